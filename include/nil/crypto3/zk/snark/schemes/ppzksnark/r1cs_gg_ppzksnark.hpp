// Forwarding header: the include path crypto3-fil-proofs names (S/core/crypto/scheme_params.hpp:29, S/core/crypto/mapped_scheme_params.hpp:32, F/caches.hpp:33) resolves to the libb2g16 shim,
// which declares the same nil::crypto3 names over the C ABI of include/b2g16.h.
#pragma once
#include <nil/crypto3/b2g16_shim.hpp>
