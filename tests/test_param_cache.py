"""Parameter-cache entry names, the .meta / .vk files and the forwarding headers
(libs/storage/include/nil/filecoin/storage/proofs/core/parameter_cache.hpp:50-94,108-145,166-173).
Host-only: runs without a GPU."""
import hashlib
import os
import subprocess

import numpy as np
import pytest

from b2g16 import lib
import util

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sha256_matches_hashlib():
    rng = np.random.Generator(np.random.PCG64(5))
    for n in list(range(0, 130)) + [255, 256, 1000, 4097]:
        data = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        assert lib.sha256(data) == hashlib.sha256(data).digest(), n
    # FIPS 180-4 known answer
    assert lib.sha256(b"abc").hex() == "ba7816bf8f01cfea414140de5dae2223b00361a396177a9cb410ff61f20015ad"


def test_cache_entry_names(monkeypatch, tmp_path):
    ident = lib.cache_identifier("stacked-proof-of-replication-merkletree-poseidon_hasher-8-0-0-sha256_hasher",
                                 "layered_drgporep::PublicParams{ graph: stacked_graph, challenges: LayerChallenges { layers: 2, max_count: 2 } }")
    prefix, digest = ident.rsplit("-", 1)
    assert prefix == "stacked-proof-of-replication-merkletree-poseidon_hasher-8-0-0-sha256_hasher"
    assert digest == hashlib.sha256(b"layered_drgporep::PublicParams{ graph: stacked_graph, challenges: LayerChallenges { layers: 2, max_count: 2 } }").hexdigest()
    monkeypatch.setenv("FIL_PROOFS_PARAMETER_CACHE", str(tmp_path) + "/")
    assert lib.cache_dir() == str(tmp_path)
    for kind, ext in ((lib.CACHE_PARAMS, "params"), (lib.CACHE_VK, "vk"), (lib.CACHE_META, "meta")):
        assert lib.cache_path(ident, kind) == "%s/v28-%s.%s" % (tmp_path, ident, ext)
    monkeypatch.delenv("FIL_PROOFS_PARAMETER_CACHE")
    assert lib.cache_path("x", lib.CACHE_VK) == "/var/tmp/filecoin-proof-parameters/v28-x.vk"
    with pytest.raises(lib.B2g16Error):
        lib.cache_path("x", 3)


def test_metadata_file(tmp_path):
    p = str(tmp_path / "v28-x.meta")
    lib.meta_write(p, 34359738368)
    assert open(p).read() == '{"sector_size":34359738368}'
    assert lib.meta_read(p) == 34359738368
    open(p, "w").write('{\n  "sector_size": 2048\n}\n')          # serde_json pretty form
    assert lib.meta_read(p) == 2048
    open(p, "w").write('{"size": 1}')
    with pytest.raises(lib.B2g16Error):
        lib.meta_read(p)
    with pytest.raises(lib.B2g16Error):
        lib.meta_read(str(tmp_path / "absent.meta"))


def test_vk_file_is_the_head_of_the_params_file(tmp_path):
    g = util.golden("groth16_m4")
    params = {n: g[n] for n in ("alpha_g1", "beta_g1", "beta_g2", "gamma_g2", "delta_g1", "delta_g2", "ic", "h", "l", "a", "b_g1", "b_g2")}
    pp, vp = str(tmp_path / "k.params"), str(tmp_path / "k.vk")
    lib.params_write(pp, params)
    lib.vk_file_write(vp, params)
    raw = open(vp, "rb").read()
    assert len(raw) == 864 + 4 + 96 * len(g["ic"])
    assert open(pp, "rb").read()[:len(raw)] == raw
    back = lib.vk_file_read(vp)
    for n in ("alpha_g1", "beta_g1", "beta_g2", "gamma_g2", "delta_g1", "delta_g2"):
        assert np.array_equal(back[n], g[n].reshape(-1)), n
    assert np.array_equal(back["ic"], g["ic"])
    open(vp, "wb").write(raw[:-1])
    with pytest.raises(lib.B2g16Error):
        lib.vk_file_read(vp)


def test_reference_include_paths_resolve(tmp_path):
    """The include lines of the reference's own headers compile unchanged against include/:
    scheme_params.hpp:29, multi_proof.hpp:32-34, caches.hpp:30-33, parameter_cache.hpp:35."""
    src = tmp_path / "inc.cpp"
    src.write_text('''
#include <nil/crypto3/algebra/curves/bls12.hpp>
#include <nil/crypto3/algebra/pairing/bls12.hpp>
#include <nil/crypto3/zk/snark/schemes/ppzksnark/r1cs_gg_ppzksnark.hpp>
#include <nil/crypto3/zk/snark/proof_systems/ppzksnark/r1cs_gg_ppzksnark.hpp>
#include <nil/crypto3/zk/snark/relations/constraint_satisfaction_problems/r1cs.hpp>
#include <nil/crypto3/algebra/multiexp.hpp>
#include <nil/crypto3/math/evaluation_domain.hpp>
using namespace nil::crypto3;
typedef zk::snark::r1cs_gg_ppzksnark<algebra::curves::bls12<381>> scheme;
int main() {
    scheme::verification_key_type vk;
    scheme::proof_type proof;
    (void)sizeof(scheme::proving_key_type);
    std::string id = b2g16::parameter_cache_identifier("prefix", "identifier");
    return (int)(vk.gamma_ABC_g1.size() + proof.bytes().size() + b2g16::parameter_cache_verifying_key_path(id).size()) == 0;
}
''')
    exe = tmp_path / "inc"
    lib_dir = os.path.join(ROOT, "crypto3-fil-proofs_b200", "lib")
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe), "-L", lib_dir, "-lb2g16",
                           "-Wl,-rpath," + lib_dir])
    subprocess.check_call([str(exe)])
