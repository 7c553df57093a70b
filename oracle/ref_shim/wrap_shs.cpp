// TEST INFRASTRUCTURE (oracle). Compiles the UNMODIFIED reference TU app/Components/Samplers/StaticHaltonSeq.cpp by textual
// inclusion so that its anonymous-namespace `randomSequence` table can be exported to the parity tests.
#include "Components/Samplers/StaticHaltonSeq.cpp"
extern "C" const float *mrt_ref_table_static_halton(unsigned *n) {
    *n = static_cast<unsigned>(randomSequence.size());
    return randomSequence.data();
}
