// TEST INFRASTRUCTURE (oracle). Compiles the UNMODIFIED reference TU app/MobileRT/Renderer.cpp by textual
// inclusion so that its anonymous-namespace `randomSequence` table can be exported to the parity tests.
#include "MobileRT/Renderer.cpp"
extern "C" const float *mrt_ref_table_tile(unsigned *n) {
    *n = static_cast<unsigned>(randomSequence.size());
    return randomSequence.data();
}
