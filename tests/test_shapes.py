"""Shape arithmetic pinned by the reference's own tests (SURVEY 8c item 4, App. G test_shapes)."""
import pytest

from b2g16 import shapes


def test_partition_count_semantics():
    # libs/storage/include/nil/filecoin/storage/proofs/core/partitions.hpp:36-38
    assert shapes.partition_count(-1) == 1
    assert shapes.partition_count(0) == -1
    assert [shapes.partition_count(p) for p in (1, 2, 10)] == [1, 2, 10]
    assert shapes.multi_proof_len(-1) == 192 and shapes.multi_proof_len(10) == 1920       # F/constants.hpp:93, F/api/seal.hpp:306
    with pytest.raises(ValueError):
        shapes.multi_proof_len(0)


def test_select_challenges_pinned_values():
    # libs/filecoin/test/parameters.cpp:35-42
    f = lambda p: shapes.select_challenges(p, 12, 11)[1]
    assert (f(1), f(2), f(4)) == (12, 6, 3)
    with pytest.raises(ValueError):
        shapes.select_challenges(0, 12, 11)


def test_porep_partition_shapes():
    # one challenge: 22 inputs including ONE (libs/storage/test/porep/stacked/circuit/proof.cpp:137-155)
    assert 1 + shapes.porep_public_inputs(1) == 22
    assert shapes.porep_partition_shape(2 << 10) == (1, 2, 39)          # BASELINE configs[1]
    assert shapes.porep_partition_shape(32 << 30) == (10, 18, 327)      # BASELINE configs[4]
    assert shapes.WINDOW_POST_CONSTRAINTS[32 << 30] <= 1 << 27           # BASELINE configs[3]: one 2^27 domain


def test_winning_post_params():
    # libs/filecoin/test/parameters.cpp:45-58
    assert shapes.winning_post_public_params(66, 1, 2048) == (66, 1, 2048)


def test_bench_uses_the_same_shapes():
    import bench
    assert bench.n_in_of("porep-c2-2KiB") == 39 and bench.n_in_of("porep-c2-32GiB") == 327
