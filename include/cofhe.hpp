// cofhe_b200 -- umbrella header of the B200 backend: put this repository's include/ BEFORE the reference's on the
// include path and `#include "cofhe.hpp"` resolves here.  It pulls in the device backend
// (cuda/cuda_cryptosystem.hpp, which replaces the reference's `#error` stub of the same path), then the reference's own
// umbrella header unchanged (#include_next), then adds what the reference's factory lacks: a make_cryptosystem that
// honours Device::GPU (reference include/cofhe.hpp:96-121 ignores the argument).
//
//   * default build: CoFHE::CPUCryptoSystem is the reference class, CoFHE::B200CryptoSystem the device backend;
//     make_cryptosystem<Device::GPU>(level, k) / make_cryptosystem(level, k, gpu_device) return the backend;
//   * -DCOFHE_B200: the name CoFHE::CPUCryptoSystem itself means the device backend, so that every factory of the
//     reference header and every driver that spells the class name (benchmarks/local.cpp, benchmarks/network.cpp,
//     examples/*.cpp) builds UNCHANGED on the B200 (tests/cpp/build.py compiles benchmarks/local.cpp this way).
#ifndef COFHE_B200_UMBRELLA_HPP_INCLUDED
#define COFHE_B200_UMBRELLA_HPP_INCLUDED

#include "cuda/cuda_cryptosystem.hpp"
#include_next "cofhe.hpp"

namespace CoFHE
{
    // tag for overload resolution: make_cryptosystem(128, 128, gpu_device{0})
    struct gpu_device
    {
        int index = 0;
    };

    inline B200CryptoSystem make_cryptosystem(uint32_t security_level, std::uint32_t k, gpu_device dev)
    {
        return B200CryptoSystem{security_level, k, false, dev.index};
    }
    inline B200CryptoSystem make_cryptosystem(SecurityLevel security_level, uint32_t k, gpu_device dev)
    {
        uint32_t bits = security_level == SecurityLevel::LOW ? 80 : security_level == SecurityLevel::MEDIUM ? 128 : 256;
        return B200CryptoSystem{bits, k, false, dev.index};
    }

    // compile-time Device dispatch (the return type depends on the device, so the enum has to be a template argument)
    template <Device D>
    inline auto make_cryptosystem(uint32_t security_level, std::uint32_t k)
    {
        if constexpr (D == Device::GPU)
            return B200CryptoSystem{security_level, k};
        else
            return CPUCryptoSystem{security_level, k};
    }
    template <Device D>
    inline auto make_cryptosystem(SecurityLevel security_level, std::uint32_t k)
    {
        uint32_t bits = security_level == SecurityLevel::LOW ? 80 : security_level == SecurityLevel::MEDIUM ? 128 : 256;
        return make_cryptosystem<D>(bits, k);
    }
} // namespace CoFHE

#endif
