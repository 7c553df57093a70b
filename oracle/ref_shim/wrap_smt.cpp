// TEST INFRASTRUCTURE (oracle). Compiles the UNMODIFIED reference TU app/Components/Samplers/StaticMersenneTwister.cpp by textual
// inclusion so that its anonymous-namespace `randomSequence` table can be exported to the parity tests.
#include "Components/Samplers/StaticMersenneTwister.cpp"
extern "C" const float *mrt_ref_table_static_mt(unsigned *n) {
    *n = static_cast<unsigned>(randomSequence.size());
    return randomSequence.data();
}
