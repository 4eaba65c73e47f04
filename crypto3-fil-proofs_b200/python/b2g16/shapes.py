"""Instance-shape arithmetic of the reference that decides what the prover is handed: how many
partitions, how many challenges (hence public inputs) per partition, how many proof bytes.
Host-side mirror, same names and error behaviour as the reference:
  partition_count    libs/storage/include/nil/filecoin/storage/proofs/core/partitions.hpp:36-38 and
                     .../core/proof/compound_proof.hpp:85-87          (-1 -> 1, 0 -> error value -1, p -> p)
  select_challenges  libs/filecoin/include/nil/filecoin/proofs/parameters.hpp:90-99, behaviour pinned by
                     libs/filecoin/test/parameters.cpp:35-42         (12 / 6 / 3 for 1 / 2 / 4 partitions)
  public-input count of one PoRep partition: generate_public_inputs,
                     libs/storage/.../porep/stacked/circuit/proof.hpp:186-269, pinned (22 with the constant
                     ONE for one challenge) by libs/storage/test/porep/stacked/circuit/proof.cpp:137-155
  proof bytes        libs/filecoin/include/nil/filecoin/proofs/constants.hpp:93 (192 per partition),
                     libs/filecoin/include/nil/filecoin/proofs/api/seal.hpp:306
"""
SINGLE_PARTITION_PROOF_LEN = 192
DRG_DEGREE = 6          # libs/storage/.../core/drgraph.hpp:49
EXP_DEGREE = 8          # libs/storage/.../porep/stacked/vanilla/graph.hpp:37
_SMALL = (2 << 10, 4 << 10, 16 << 10, 32 << 10, 8 << 20, 16 << 20, 512 << 20, 1 << 30)
_LARGE = (32 << 30, 64 << 30)
POREP_MINIMUM_CHALLENGES = {**{k: 2 for k in _SMALL}, **{k: 176 for k in _LARGE}}     # F/constants.hpp:65-68
POREP_PARTITIONS = {**{k: 1 for k in _SMALL}, **{k: 10 for k in _LARGE}}              # F/constants.hpp:70-73
LAYERS = {**{k: 2 for k in _SMALL}, **{k: 11 for k in _LARGE}}                        # F/constants.hpp:75-78
WINDOW_POST_SECTOR_COUNT = {**{k: 2 for k in _SMALL}, 32 << 30: 2349, 64 << 30: 2300}  # F/constants.hpp:85-90
WINDOW_POST_CONSTRAINTS = {32 << 30: 125279217, 64 << 30: 129887900}                  # comments at F/constants.hpp:88-89


def partition_count(partitions):
    """-1 (none requested) -> 1; 0 -> -1, the reference's error value; p -> p."""
    return 1 if partitions == -1 else (-1 if partitions == 0 else partitions)


def select_challenges(partitions, minimum_total_challenges, layers):
    """Smallest per-partition challenge count whose total over all partitions reaches the minimum.
    Returns (layers, count) -- LayerChallenges; challenges_count_all() is `count`."""
    if partitions <= 0:
        raise ValueError("select_challenges: partitions must be positive")
    count = 1
    while partitions * count < minimum_total_challenges:
        count += 1
    return layers, count


def porep_public_inputs(challenges):
    """Public inputs of one stacked-PoRep partition circuit, excluding the constant ONE: replica_id, comm_d,
    comm_r, then per challenge one comm_d inclusion, DRG_DEGREE + EXP_DEGREE parents, the challenge index and
    two more inclusions (= 18)."""
    return 3 + (1 + DRG_DEGREE + EXP_DEGREE + 1 + 2) * challenges


def porep_partition_shape(sector_bytes):
    """(partitions, challenges per partition, public inputs per partition) for a registered sector size."""
    p = POREP_PARTITIONS[sector_bytes]
    _, ch = select_challenges(p, POREP_MINIMUM_CHALLENGES[sector_bytes], LAYERS[sector_bytes])
    return p, ch, porep_public_inputs(ch)


def multi_proof_len(partitions):
    n = partition_count(partitions)
    if n < 0:
        raise ValueError("multi_proof_len: zero partitions")
    return SINGLE_PARTITION_PROOF_LEN * n


def winning_post_public_params(challenge_count, sector_count, sector_size):
    """libs/filecoin/test/parameters.cpp:45-58: Winning PoSt swaps the two counts (66 challenges over 1 sector
    become 66 'sectors' with 1 challenge each).  Returns (sector_count, challenge_count, sector_size)."""
    return challenge_count, sector_count, sector_size
