"""A model of the DcpExchange step protocol (vllm_metax_b200/attention/dcp.py, csrc/dcp_exchange.cu) on plain threads.

The GPU tests show that the fused decode-context-parallel step gives the right numbers; this test is about WHY one set
of buffers per rank is enough now that the step number lives in device memory (round 1 alternated two sets by step
parity).  Every rank is a thread that runs the three launches of a step in stream order

    gather(k):    push my query slice into every rank's gathered query, publish gather flag k, wait for all flags >= k
    attention(k): read my gathered query, write my partial (out, lse) buffer
    fold(k):      publish fold flag k, wait for all flags >= k, read every rank's partial buffer

against shared "buffers" that only remember which step last wrote them.  A reader checks the version when it starts and
again when it ends, with a random pause in between, so a writer that gets in too early is caught.  Ranks are skewed by
random sleeps at every point.  The negative control removes the wait of the gather launch — the one wait the argument
rests on — and the model must then catch the overwrite.
"""
import random
import threading
import time

import pytest


class Model:
    def __init__(self, world, gather_waits=True, seed=0):
        self.n = world
        self.gather_waits = gather_waits
        self.q = [[0] * world for _ in range(world)]      # q[dst][src]: step whose slice of src lies in dst's gathered query
        self.out = [0] * world                            # out[r]: step whose partial lies in rank r's buffer
        self.gflag = [[0] * world for _ in range(world)]  # gflag[dst][src]
        self.fflag = [[0] * world for _ in range(world)]
        self.errors = []
        self.rnd = [random.Random(seed * 100 + r) for r in range(world)]
        self.slow = {}                                    # (rank, phase, step) -> extra seconds (adversarial schedules)

    def pause(self, r, phase, k, scale=0.0015):
        time.sleep(self.rnd[r].random() * scale + self.slow.get((r, phase, k), 0.0))

    def wait(self, flags, k, deadline):
        while any(f < k for f in flags):
            if time.time() > deadline:
                raise TimeoutError("a rank never arrived")
            time.sleep(0.0001)

    def rank(self, r, steps):
        n = self.n
        deadline = time.time() + 60
        try:
            for k in range(1, steps + 1):
                # ---- gather launch ----
                self.pause(r, "gather", k)
                for dst in range(n):
                    self.q[dst][r] = k
                for dst in range(n):
                    self.gflag[dst][r] = k
                if self.gather_waits:
                    self.wait(self.gflag[r], k, deadline)
                # ---- attention launch: reads the gathered query, writes the partial ----
                if self.gather_waits:
                    seen = list(self.q[r])
                    if any(v != k for v in seen):
                        self.errors.append("rank %d step %d: gathered query holds steps %s at the start" % (r, k, seen))
                self.pause(r, "attention", k)
                if self.gather_waits:
                    seen = list(self.q[r])
                    if any(v != k for v in seen):
                        self.errors.append("rank %d step %d: gathered query overwritten while read: %s" % (r, k, seen))
                self.out[r] = k
                # ---- fold launch ----
                self.pause(r, "fold", k)
                for dst in range(n):
                    self.fflag[dst][r] = k
                self.wait(self.fflag[r], k, deadline)
                seen = list(self.out)
                if any(v != k for v in seen):
                    self.errors.append("rank %d step %d: partials hold steps %s at the start of the fold" % (r, k, seen))
                self.pause(r, "fold_read", k)
                seen = list(self.out)
                if any(v != k for v in seen):
                    self.errors.append("rank %d step %d: a partial was overwritten while the fold read it: %s" % (r, k, seen))
        except TimeoutError as e:
            self.errors.append("rank %d: %s" % (r, e))

    def run(self, steps):
        ts = [threading.Thread(target=self.rank, args=(r, steps)) for r in range(self.n)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        return self.errors


@pytest.mark.parametrize("world", [2, 4, 8])
def test_one_buffer_set_is_enough_with_a_gather_in_every_step(world):
    m = Model(world, seed=world)
    # on top of the random skew: one rank that dawdles inside its fold read, one that is late to every gather
    for k in range(1, 40, 5):
        m.slow[(0, "fold_read", k)] = 0.01
        m.slow[(world - 1, "gather", k + 1)] = 0.01
    assert m.run(40) == []


def test_the_model_catches_a_protocol_without_the_gather_wait():
    """Negative control: with the gather launch returning at once, a fast rank's next attention launch overwrites its
    partial while a slow rank's fold is still reading it — which is what the second buffer set of round 1 was for."""
    m = Model(2, gather_waits=False, seed=1)
    m.slow[(0, "fold_read", 1)] = 0.05
    errs = m.run(3)
    assert any("overwritten while the fold read it" in e or "at the start of the fold" in e for e in errs), errs
