"""A timing model of the pipelined sample reduce (bench.py ReducePipeline, include/rt_multi.h rt_peer_reduce):
every rank renders frame k on its main stream while its reduce of frame k-1 runs on a side stream. The model
replays the orderings the code enforces — stream order, the events, the entry wait of the reduce kernel (it
proceeds once EVERY rank has started the reduce of that epoch) — with random durations, and checks the one
hazard the buffer rotation exists for: no rank may render into a partial buffer while any rank still reads it.
Three buffers (sharding.PeerReducer.NBUF) are safe under every schedule; the same orderings with two are not.
CPU only: no GPU, no process group."""
import random

from opengl_raytracing_b200 import sharding


def simulate(world: int, frames: int, nbuf: int, rng: random.Random):
    """Returns the list of hazards (writer rank, frame, reader rank, epoch) of one random schedule."""
    render_end = [[0.0] * (frames + 1) for _ in range(world)]
    render_iv = [[None] * (frames + 1) for _ in range(world)]
    red_start = [[0.0] * (frames + 1) for _ in range(world)]
    red_end = [[0.0] * (frames + 1) for _ in range(world)]
    read_iv = [[None] * (frames + 1) for _ in range(world)]
    speed = [rng.uniform(0.5, 2.0) for _ in range(world)]          # some ranks are simply slower
    for k in range(1, frames + 1):
        for r in range(world):
            # render k: behind render k-1 (stream order) and behind this rank's reduce of frame k-2 (the event)
            t = render_end[r][k - 1]
            if k >= 3:
                t = max(t, red_end[r][k - 2])
            d = rng.uniform(0.2, 1.0) * speed[r]
            render_iv[r][k] = (t, t + d)
            render_end[r][k] = t + d
        for r in range(world):
            # reduce k starts behind render k (event) and reduce k-1 (stream order); at its start it raises its flag
            red_start[r][k] = max(render_end[r][k], red_end[r][k - 1])
        passed = max(red_start[r][k] for r in range(world))        # every flag of epoch k is up
        for r in range(world):
            d = rng.uniform(0.05, 3.0)                              # from far shorter to far longer than a render
            read_iv[r][k] = (passed, passed + d)
            red_end[r][k] = passed + d
    hazards = []
    for a in range(world):
        for j in range(1, frames + 1):
            w0, w1 = render_iv[a][j]
            for k in range(1, j):
                if k % nbuf != j % nbuf:
                    continue
                for b in range(world):                              # rank b's reduce of epoch k reads rank a's buffer k % nbuf
                    r0, r1 = read_iv[b][k]
                    if w0 < r1 and r0 < w1:
                        hazards.append((a, j, b, k))
    return hazards


def test_three_rotating_partials_are_safe_under_every_schedule():
    assert sharding.PeerReducer.NBUF == 3
    rng = random.Random(1234)
    for trial in range(400):
        world = rng.choice([2, 3, 4, 8])
        assert simulate(world, 12, sharding.PeerReducer.NBUF, rng) == [], trial


def test_two_partials_would_race_when_the_reduce_overlaps_the_next_render():
    rng = random.Random(99)
    found = sum(1 for _ in range(200) if simulate(rng.choice([2, 4, 8]), 12, 2, rng))
    assert found > 0          # the third buffer is not decoration


def test_render_target_follows_the_epoch():
    """PeerReducer.partial is the buffer of the NEXT reduce: (epoch + 1) % NBUF; reduce() reads epoch % NBUF."""
    n = sharding.PeerReducer.NBUF
    epoch, seen = 0, []
    for _ in range(7):
        target = (epoch + 1) % n          # what .partial returns before the reduce is enqueued
        epoch += 1                        # reduce(): self.epoch += 1, reads partial[self.epoch % NBUF]
        assert epoch % n == target
        seen.append(target)
    assert seen[:n] != [seen[0]] * n and len(set(seen)) == n
