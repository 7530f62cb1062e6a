"""MapPoint::ComputeDistinctiveDescriptors — reference src/MapPoint.cc:246-312: of the descriptors under which a
map point was observed, the one with the least median Hamming distance to the rest.  CPU: known answers + a numpy restatement.
GPU: batches of map points through the C ABI against the oracle, bit-exact (index and median)."""
import numpy as np
import pytest

from oracle import oracle_lib as O


def numpy_restatement(desc):
    n = desc.shape[0]
    if n == 0:
        return -1, -1
    d = np.unpackbits(desc[:, None, :] ^ desc[None, :, :], axis=2).sum(axis=2)
    med = np.sort(d, axis=1)[:, int(0.5 * (n - 1))]
    i = int(np.argmin(med))          # first index of the minimum = strict '<' of :300
    return i, int(med[i])


def groups(seed, n_points, max_obs):
    rng = np.random.default_rng(seed)
    sizes = rng.integers(0, max_obs + 1, n_points)
    if n_points > 8:
        sizes[:4] = [0, 1, 2, 3]
    else:
        sizes[:] = np.maximum(sizes, max_obs // 2)
    start = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    centre = rng.integers(0, 256, (n_points, 32), dtype=np.uint8)
    desc = np.repeat(centre, sizes, axis=0)
    noise = rng.integers(0, 256, desc.shape, dtype=np.uint8) & rng.integers(0, 256, desc.shape, dtype=np.uint8) & \
        rng.integers(0, 256, desc.shape, dtype=np.uint8)
    desc ^= noise * (rng.random((desc.shape[0], 1)) < 0.8)   # some exact duplicates -> ties
    return start, desc


def test_oracle_kat_and_numpy():
    z = np.zeros((1, 32), np.uint8)
    assert O.distinctive_descriptors([0, 1], z)[0][0] == 0 and O.distinctive_descriptors([0, 1], z)[1][0] == 0
    assert list(O.distinctive_descriptors([0, 0], np.zeros((0, 32), np.uint8))[0]) == [-1]
    # three descriptors on a line 0 -- 8 bits -- 24 bits: rows sorted (0,8,24), (0,8,16), (0,16,24); median index 1 -> 8, 8, 16
    d = np.zeros((3, 32), np.uint8)
    d[1, 0] = 0xFF
    d[2, 0:3] = 0xFF
    idx, med = O.distinctive_descriptors([0, 3], d)
    assert (idx[0], med[0]) == (0, 8)            # tie between rows 0 and 1: the first wins
    # two descriptors: median index int(0.5 * 1) = 0 -> the self distance 0, first wins
    assert tuple(x[0] for x in O.distinctive_descriptors([0, 2], d[:2])) == (0, 0)
    start, desc = groups(3, 60, 25)
    idx, med = O.distinctive_descriptors(start, desc)
    for p in range(60):
        assert (idx[p], med[p]) == numpy_restatement(desc[start[p]:start[p + 1]])


@pytest.mark.gpu
@pytest.mark.parametrize("seed,n_points,max_obs", [(1, 500, 12), (2, 64, 90), (3, 3, 300), (4, 2000, 6)])
def test_distinctive_descriptors_gpu(seed, n_points, max_obs):
    import orb_slam_cuda_b200 as X
    start, desc = groups(seed, n_points, max_obs)
    r_idx, r_med = O.distinctive_descriptors(start, desc)
    idx, med = X.ComputeDistinctiveDescriptors(start, desc)
    assert np.array_equal(idx, r_idx) and np.array_equal(med, r_med)
    assert (r_med[start[1:] - start[:-1] >= 3] > 0).any()
    with pytest.raises(X.OrbxError):
        X.ComputeDistinctiveDescriptors(np.array([1, 2], np.int32), desc)
