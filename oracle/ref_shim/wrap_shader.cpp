// TEST INFRASTRUCTURE (oracle). Compiles the UNMODIFIED reference TU app/MobileRT/Shader.cpp by textual
// inclusion so that its anonymous-namespace `randomSequence` table can be exported to the parity tests.
#include "MobileRT/Shader.cpp"
extern "C" const float *mrt_ref_table_shader(unsigned *n) {
    *n = static_cast<unsigned>(randomSequence.size());
    return randomSequence.data();
}

// Shader::getLightIndex keeps its draw counter in a function-local static that lives for the whole process and is never
// reset (Shader.cpp:224-225). The replay tests need its current value; the local symbol is reached by its assembler name.
extern ::std::atomic<::std::uint32_t> mrt_ref_light_index_sampler __asm__("_ZZN8MobileRT6Shader13getLightIndexEvE7sampler");
extern "C" unsigned mrt_ref_light_index_counter() { return mrt_ref_light_index_sampler.load(); }
