// Forwarding header: the include path crypto3-fil-proofs names (S/core/proof/multi_proof.hpp:32, F/caches.hpp:30, S/core/parameter_cache.hpp:35) resolves to the libb2g16 shim,
// which declares the same nil::crypto3 names over the C ABI of include/b2g16.h.
#pragma once
#include <nil/crypto3/b2g16_shim.hpp>
