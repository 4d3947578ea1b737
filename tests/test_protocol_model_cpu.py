"""A schedule-exploring model of the device-side backup-worker protocol (csrc/comm.cu: backup_arrive_and_select,
backup_finish, block_barrier), the race-detection tier of SURVEY 5.2 that needs no GPU.

Every rank is a coroutine that yields before each access to memory another rank can see (ticket counters, arrival
words, barrier flags); a seeded scheduler interleaves them arbitrarily, including long stalls of a "straggler". Checked
for every schedule: no deadlock, every owner selects exactly R distinct ranks per call, the selected ranks are the
first R to have taken that owner's tickets of the call, and no owner ever counts a ticket of another call. Two mutants
-- the protocol without its exit barrier, and with a call counter that advances at a different pace on one rank (the
per-block-counter bug fixed in round 2) -- must be caught by the same checks, which shows the model can fail."""
import random

import pytest

U32 = 1 << 32


class World(object):
    def __init__(self, n):
        self.n = n
        self.ticket = [0] * n                       # per owner: TICKET_OFF
        self.arrive = [[0] * n for _ in range(n)]   # per owner, per rank: ARRIVE_OFF + rank
        self.call = [0] * n                         # per rank: BK_CALL_OFF
        self.flag = [[0] * n for _ in range(n)]     # barrier page of rank p: epoch published by rank r
        self.epoch = [0] * n
        self.order = [[[] for _ in range(n)] for _ in range(64)]  # ground truth: arrival order per (call, owner)
        self.selected = {}                          # (call, owner) -> tuple of ranks


def barrier(w, r):
    w.epoch[r] += 1
    e = w.epoch[r]
    for p in range(w.n):
        yield
        w.flag[p][r] = e
    while True:
        yield
        if all(w.flag[r][p] >= e for p in range(w.n)):
            return


def rank_program(w, r, R, calls, exit_barrier=True, call_stride=1):
    for true_call in range(calls):
        yield
        k = w.call[r]
        for o in range(w.n):                        # block 0: one thread per owner takes a ticket there
            yield
            t = w.ticket[o]
            w.ticket[o] = (t + 1) % U32             # atomicAdd_system
            w.order[true_call][o].append(r)
            yield
            w.arrive[o][r] = (t + 1) % U32          # st.release
        lo = (k * w.n) % U32
        while True:                                 # owner duty: wait for the first R arrivals of call k
            yield
            mine = []
            for q in range(w.n):
                a = w.arrive[r][q]
                ordv = (a - 1 - lo) % U32           # wraps to a huge value for tickets of earlier calls
                if a != 0 and ordv < R:
                    mine.append(q)
            if len(mine) >= R:
                break
        w.selected[(true_call, r)] = tuple(mine)
        if exit_barrier:
            yield from barrier(w, r)
        yield
        w.call[r] = k + call_stride                 # backup_finish (last block of the call)


def run(n, R, calls, seed, straggler=None, **mutant):
    w = World(n)
    rng = random.Random(seed)
    progs = {r: rank_program(w, r, R, calls, **({} if r != mutant.get("rank", r) else
                                                 {k: v for k, v in mutant.items() if k != "rank"})) for r in range(n)}
    steps = 0
    while progs:
        steps += 1
        if steps > 400000:
            return w, "deadlock"
        live = list(progs)
        if straggler in progs and len(live) > 1 and rng.random() < 0.97:
            live.remove(straggler)                  # the straggler almost never gets the GPU
        r = rng.choice(live)
        try:
            next(progs[r])
        except StopIteration:
            del progs[r]
    return w, "done"


def check(w, n, R, calls):
    for c in range(calls):
        for o in range(n):
            sel = w.selected.get((c, o))
            assert sel is not None, ("owner never selected", c, o)
            assert len(sel) == R and len(set(sel)) == R, ("wrong count", c, o, sel)
            assert set(sel) == set(w.order[c][o][:R]), ("not the first R arrivals of this call", c, o, sel,
                                                        w.order[c][o])


@pytest.mark.parametrize("n,R", [(2, 1), (3, 2), (4, 2), (4, 3), (8, 6)])
def test_backup_worker_protocol_for_random_schedules(n, R):
    for seed in range(40 if n < 8 else 12):
        straggler = None if seed % 3 else seed % n
        w, status = run(n, R, calls=5, seed=seed, straggler=straggler)
        assert status == "done", ("deadlock", n, R, seed)
        check(w, n, R, 5)
        assert all(c == 5 for c in w.call) and all(t == 5 * n for t in w.ticket)


def _mutant_is_caught(**mutant):
    for seed in range(60):
        w, status = run(4, 2, calls=5, seed=seed, straggler=seed % 4, **mutant)
        if status != "done":
            return True
        try:
            check(w, 4, 2, 5)
        except AssertionError:
            return True
    return False


def test_the_model_catches_a_protocol_without_exit_barrier():
    # without the barrier a fast rank takes tickets of call k+1 while a slow owner still selects for call k
    assert _mutant_is_caught(exit_barrier=False)


def test_the_model_catches_a_call_counter_out_of_step():
    # one rank whose call counter advances by two per call (the per-block counter bug): its window never matches
    assert _mutant_is_caught(rank=1, call_stride=2)
