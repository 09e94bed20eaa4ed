"""The batched GPU LLL (bltcu_lll_batch) against the oracle's exact LLL (oracle/lll.py, the restatement of
NTL::LLL(det2, B, U), libblt/ntl_wrapper.cc:37-52) and against the host implementation (bltcu_lll): same rank, same
reduced basis, same transform for every problem."""
import numpy as np
import pytest

from oracle.lll import lll as oracle_lll

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    import blt_b200
    return blt_b200


def _compare(gpu, Bs, exact_u=True, oracle_first=10**9):
    ranks, Rs, Us, info = gpu.lll_batch(Bs)
    for i, B in enumerate(Bs):
        rk, Br, U = gpu.lll(B)
        assert ranks[i] == rk
        assert (Rs[i] == Br).all()
        assert (Us[i] == U).all()
        if i < oracle_first:  # the CPU oracle itself, not only the product's host routine
            ork, oB, oU = oracle_lll([[int(x) for x in row] for row in B])
            wrap = lambda M: np.array([[((int(x) + 2**63) % 2**64) - 2**63 for x in row] for row in M], dtype=np.int64)
            assert ranks[i] == ork and (Rs[i] == wrap(oB)).all() and (Us[i] == wrap(oU)).all()
        if exact_u:  # U * B_in = B_out in exact integers (when the entries of U fit int64, see below)
            assert (Us[i].astype(object) @ B.astype(object) == Rs[i].astype(object)).all()
    return info


def test_config5_batch(gpu):
    """A in Z^{20x20} uniform on [-100, 100]; the lattice basis is A^T."""
    rng = np.random.default_rng(5)
    Bs = np.stack([rng.integers(-100, 101, size=(20, 20)).T for _ in range(300)])
    info = _compare(gpu, Bs)
    # the fp64-decision pass completes all but a handful (decisions on a tie); a handful goes to the host routine on all
    # cores, a crowd would go to the 512-bit kernel
    assert info["deferred"] <= 15 and info["on_host"] == info["deferred"]


def test_exact_kernel_alone(gpu, monkeypatch):
    """BLTCU_LLL_EXACT_ONLY=1: every problem through the 512-bit kernel (the round-1 path), same output."""
    monkeypatch.setenv("BLTCU_LLL_EXACT_ONLY", "1")
    rng = np.random.default_rng(6)
    Bs = np.stack([rng.integers(-100, 101, size=(16, 16)) for _ in range(40)])
    info = _compare(gpu, Bs)
    assert info["deferred"] == 40 and info["on_host"] == 0


def test_config2_dependent_rows(gpu):
    """Config 2: 40 variables, 20 rows => a 40 x 20 generating set of rank 20 (the dependent-row branch of the LLL)."""
    rng = np.random.default_rng(2)
    Bs = np.stack([rng.integers(-100, 101, size=(40, 20)) for _ in range(12)])
    ranks, Rs, Us, info = gpu.lll_batch(Bs)
    assert (ranks == 20).all()
    for i in range(len(Bs)):
        assert not Rs[i][:20].any()  # zero rows first, like NTL leaves them
    # the transform of a dependent generating set has entries beyond int64; both implementations return their low 64
    # bits, as the reference does (to_long, libblt/ntl_wrapper.cc:159-168), so only equality is checked here
    _compare(gpu, Bs[:6], exact_u=False)


def test_overflow_falls_back_to_the_host_implementation(gpu):
    rng = np.random.default_rng(9)
    small = [rng.integers(-50, 51, size=(12, 12)) for _ in range(6)]
    huge = [rng.integers(-10**17, 10**17, size=(12, 12)) for _ in range(3)]  # Gram determinants beyond 512 bits
    Bs = np.stack([small[0], huge[0], small[1], small[2], huge[1], small[3], huge[2], small[4], small[5]])
    info = _compare(gpu, Bs, exact_u=False)
    assert info["on_host"] == 3


def test_edge_shapes(gpu):
    ranks, Rs, Us, info = gpu.lll_batch(np.zeros((0, 3, 3), dtype=np.int64))
    assert len(ranks) == 0
    _compare(gpu, np.array([[[5]], [[0]], [[-7]]]))
    _compare(gpu, np.array([[[1, 2, 3]], [[0, 0, 0]]]))
    _compare(gpu, np.array([[[2, 0], [0, 2], [1, 1]]]))
