"""Model check of the warp-cooperative rejection tail (csrc/device_core.cuh coop_ball_tail, used by csrc/scan.cu):
the lane/segment arithmetic — candidates per failing lane = largest power of two with failing * per <= 32, lane L
serves failing lane L >> shift with candidate kbase + (L & (per - 1)), a failing lane takes the lowest accepted lane
of its segment — must hand every lane the FIRST accepted k >= 1 of its own stream, i.e. what the sequential loop of
sample.rs:44-52 returns.  The accept pattern of each lane's stream is drawn at random; the GPU tests check the real
kernel against the oracle."""
import numpy as np


def clz32(x):
    return 32 - int(x).bit_length()


def cooperative(accept, failed):
    """accept[lane][k] -> bool for k >= 1; failed: set of lanes whose candidate 0 was rejected. Returns {lane: k}."""
    failed = set(failed)
    out = {}
    kbase = 1
    while failed:
        order = sorted(failed)
        nf = len(order)
        shift = 5 - (0 if nf <= 1 else 32 - clz32(nf - 1))
        per = 1 << shift
        assert nf * per <= 32 and (per == 32 or nf * per * 2 > 32 or nf == 1)
        ok = [False] * 32
        for lane in range(32):  # every lane evaluates one candidate
            j = lane >> shift
            if j < nf:
                ok[lane] = accept[order[j]][kbase + (lane & (per - 1))]
        for my_j, lane in enumerate(order):
            seg = [ok[(my_j << shift) + i] for i in range(per)]
            if any(seg):
                out[lane] = kbase + seg.index(True)  # __ffs: lowest lane of the segment = lowest k
                failed.discard(lane)
        kbase += per
    return out


def test_cooperative_tail_returns_the_first_accepted_candidate():
    rng = np.random.default_rng(7)
    for trial in range(400):
        p_accept = rng.choice([0.05, 0.3, 0.524, 0.9])
        accept = rng.random((32, 4096)) < p_accept
        accept[:, 4000:] = True  # every stream accepts eventually
        n_failed = int(rng.integers(0, 33))
        failed = rng.choice(32, size=n_failed, replace=False).tolist()
        got = cooperative(accept, failed)
        assert set(got) == set(failed)
        for lane in failed:
            first = 1 + int(np.argmax(accept[lane, 1:]))
            assert got[lane] == first, (trial, lane, got[lane], first)


def test_candidates_per_failing_lane():
    for nf, per in {1: 32, 2: 16, 3: 8, 4: 8, 5: 4, 8: 4, 9: 2, 16: 2, 17: 1, 32: 1}.items():
        shift = 5 - (0 if nf <= 1 else 32 - clz32(nf - 1))
        assert (1 << shift) == per
