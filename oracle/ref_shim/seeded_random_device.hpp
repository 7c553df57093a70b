// TEST INFRASTRUCTURE (oracle). Force-included (-include) into every reference TU when the
// unmodified reference sources under /root/reference are compiled into oracle/_ref/libmobilert_ref.so.
// It swaps std::random_device for a seedable stub so that the reference's table shuffles
// (app/MobileRT/Utils/Utils.hpp:215,232,248; Components/Samplers/MersenneTwister.cpp:7; PCG.cpp:7)
// are reproducible run to run. The reference sources stay byte-identical.
#ifndef MRT_ORACLE_SEEDED_RANDOM_DEVICE_HPP
#define MRT_ORACLE_SEEDED_RANDOM_DEVICE_HPP
#ifdef __cplusplus
#include <random>
#include <cstdlib>
#include <string>
namespace std {
    class mrt_seeded_random_device {
    public:
        using result_type = unsigned int;
        mrt_seeded_random_device() {}
        explicit mrt_seeded_random_device(const std::string &) {}
        mrt_seeded_random_device(const mrt_seeded_random_device &) = delete;
        static constexpr result_type min() { return 0u; }
        static constexpr result_type max() { return 0xFFFFFFFFu; }
        double entropy() const noexcept { return 0.0; }
        // draws made so far by every instance (the shim can rewind it: ref_random_device_counter in ref_api.cpp)
        static unsigned int &counter() { static unsigned int c = 0u; return c; }
        result_type operator()() {
            // Every draw returns seed + draw-counter, seed from MRT_ORACLE_SEED (default 1).
            const char *s = ::getenv("MRT_ORACLE_SEED");
            const unsigned int seed = s ? static_cast<unsigned int>(::strtoul(s, nullptr, 10)) : 1u;
            return seed + 7919u * (counter()++);
        }
    };
}
#define random_device mrt_seeded_random_device
#endif
#endif
