"""The exchange cycle of `DistRuntime.run_resident_cycle` (parallel/dist.py) as a host model.

Halo rows are exchanged every `cycle` passes only; the passes between compute a shrinking margin of
ghost rows themselves, and the cycles alternate between two pairs of band buffers so that no "I am
done reading your rows" flag is needed. This test replays that schedule with NumPy band buffers and
the C++ host filters, with the ranks interleaved at random (a rank may run as far ahead of its
neighbours as its one wait per cycle lets it, and pushes land in the neighbour's memory the moment
they are issued), and compares the gathered image with the filter applied to the whole image. A
write-after-read hazard in the buffer rotation, a ghost row used before it is valid or a wrong launch
extent shows up as a mismatch under some interleaving.

The GPU version is checked bit for bit in scripts/dist_check.py and inside every `bench.py --gpus N`
run; the reference has no halo exchange at all (kernel.cu:195 filters bands in isolation).
"""
import numpy as np
import pytest

import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.chain import band_rows

G = 8  # ghost rows above and below a resident band (parallel/dist.py)


class Rank:
    def __init__(self, r, world, img, rng):
        H = img.shape[0]
        self.r, self.world, self.H = r, world, H
        self.y0, self.y1 = band_rows(H, world, r)
        rows = self.y1 - self.y0
        # four band buffers full of noise: a row that is used before it is written poisons the result
        self.buf = [rng.integers(0, 256, (rows + 2 * G,) + img.shape[1:], dtype=np.uint8) for _ in range(4)]
        a, b = max(0, self.y0 - G), min(H, self.y1 + G)
        self.buf[0][a - (self.y0 - G):b - (self.y0 - G)] = img[a:b]   # the scatter: band + ghost rows
        self.flag_up = self.flag_dn = 0    # epochs the neighbours have signalled
        self.p = 0                         # next pass

    def rows(self, k, a, b):
        return self.buf[k][a - (self.y0 - G):b - (self.y0 - G)]


def run_model(chain, halo, img, world, passes, cycle, rng, order):
    ranks = [Rank(r, world, img, rng) for r in range(world)]
    H = img.shape[0]

    def ready(k):
        if k.p >= passes:
            return False
        if k.p % cycle == 0 and k.p > 0:   # the one wait of a cycle: the rows the neighbours pushed
            if k.r > 0 and k.flag_up < k.p:
                return False
            if k.r + 1 < world and k.flag_dn < k.p:
                return False
        return True

    def step(k):
        p = k.p
        c, j = (p // cycle) & 1, p % cycle
        pair, other = ((0, 1), 2) if c == 0 else ((2, 3), 0)
        last = j == cycle - 1
        inb = pair[j & 1]
        outb = other if last else pair[(j & 1) ^ 1]
        ext = 0 if last else (cycle - 1 - j) * halo
        ya = k.y0 - (ext if k.r > 0 else 0)
        yb = k.y1 + (ext if k.r + 1 < world else 0)
        a, b = max(0, ya - halo), min(H, yb + halo)             # input rows of the launch
        out = b2.host_run(chain, np.ascontiguousarray(k.rows(inb, a, b)))
        k.rows(outb, ya, yb)[:] = out[ya - a:yb - a]
        if last and p + 1 < passes:                              # push cycle * halo rows, then signal
            d = cycle * halo
            if k.r > 0:
                up = ranks[k.r - 1]
                up.rows(outb, k.y0, k.y0 + d)[:] = k.rows(outb, k.y0, k.y0 + d)
                up.flag_dn = p + 1
            if k.r + 1 < world:
                dn = ranks[k.r + 1]
                dn.rows(outb, k.y1 - d, k.y1)[:] = k.rows(outb, k.y1 - d, k.y1)
                dn.flag_up = p + 1
        k.p += 1

    while any(k.p < passes for k in ranks):
        can = [k for k in ranks if ready(k)]
        assert can, "deadlock: no rank can run"
        if order == "random":
            step(can[int(rng.integers(len(can)))])
        elif order == "ahead":       # the lowest rank that can run goes as far as it can
            step(can[0])
        else:                        # "behind": the highest rank first
            step(can[-1])
    final = 2 if (passes // cycle) % 2 == 1 else 0
    return np.concatenate([k.rows(final, k.y0, k.y1) for k in ranks])


@pytest.mark.parametrize("chain,halo", [("gauss5", 2), ("gauss3", 1), ("gauss5@replicate", 2)])
@pytest.mark.parametrize("cycle", [2, 4])
@pytest.mark.parametrize("world,H", [(2, 40), (3, 61), (4, 64)])
def test_exchange_cycle_schedule_is_exact_under_any_interleaving(chain, halo, cycle, world, H):
    if cycle * halo > G:
        pytest.skip("more ghost rows than the band buffers carry")
    W, passes = 24, 2 * cycle * 2
    img = b2.synth_image(W, H, 3, seed=cycle + world)
    want = img
    for _ in range(passes):
        want = b2.host_run(chain, want)
    for seed, order in enumerate(("ahead", "behind", "random", "random", "random")):
        rng = np.random.default_rng(1000 * cycle + 10 * world + seed)
        got = run_model(chain, halo, img, world, passes, cycle, rng, order)
        assert np.array_equal(got, want), f"{chain} cycle {cycle} world {world} order {order}"


def test_a_single_pair_of_buffers_would_race():
    """The reason for the two pairs: with one pair the pushing pass would overwrite ghost rows a slower
    neighbour still has to read. The model shows it (the same schedule, pushes aimed at the pair's own
    first buffer, the fast rank first)."""
    chain, halo, cycle, world, H, W = "gauss5", 2, 2, 2, 40, 24
    img = b2.synth_image(W, H, 3, seed=9)
    passes = 4
    want = img
    for _ in range(passes):
        want = b2.host_run(chain, want)
    rng = np.random.default_rng(1)
    ranks = [Rank(r, world, img, rng) for r in range(world)]

    def step(k):
        p = k.p
        j = p % cycle
        last = j == cycle - 1
        inb, outb = j & 1, (j & 1) ^ 1                      # one pair only
        ext = 0 if last else (cycle - 1 - j) * halo
        ya = k.y0 - (ext if k.r > 0 else 0)
        yb = k.y1 + (ext if k.r + 1 < world else 0)
        a, b = max(0, ya - halo), min(H, yb + halo)
        out = b2.host_run(chain, np.ascontiguousarray(k.rows(inb, a, b)))
        k.rows(outb, ya, yb)[:] = out[ya - a:yb - a]
        if last and p + 1 < passes:
            d = cycle * halo
            for nb, lo, hi in ((k.r - 1, k.y0, k.y0 + d), (k.r + 1, k.y1 - d, k.y1)):
                if 0 <= nb < world:
                    ranks[nb].rows(outb, lo, hi)[:] = k.rows(outb, lo, hi)
        k.p += 1

    # rank 0 runs its whole first cycle (and pushes into rank 1's buffer 0) before rank 1 has started:
    # rank 1's first pass then reads ghost rows of the wrong pass
    step(ranks[0]); step(ranks[0])
    step(ranks[1]); step(ranks[1])
    step(ranks[0]); step(ranks[1]); step(ranks[0]); step(ranks[1])
    got = np.concatenate([k.rows(0, k.y0, k.y1) for k in ranks])
    assert not np.array_equal(got, want)
