// TEST INFRASTRUCTURE (oracle). Compiles the UNMODIFIED reference TU app/MobileRT/Shader.cpp by textual
// inclusion so that its anonymous-namespace `randomSequence` table can be exported to the parity tests.
#include "MobileRT/Shader.cpp"
extern "C" const float *mrt_ref_table_shader(unsigned *n) {
    *n = static_cast<unsigned>(randomSequence.size());
    return randomSequence.data();
}
