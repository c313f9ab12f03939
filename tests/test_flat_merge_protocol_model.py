"""A model of how the flat schedule finishes a (sequence, kv head) pair that several CTAs worked on
(vllm_metax_b200/csrc/attention_stream.cu, the merge at the end of a segment).

Every piece of the pair is a thread.  A piece reads the pair's arrival counter when its segment begins and once more
shortly before it ends; at the end it either
  * found every OTHER piece delivered -> merges straight from its own result plus the others' published partials,
    writes the output and zeroes the counter ("direct" path, no partial written, counter not incremented), or
  * publishes its partial, increments the counter, and merges from the published partials if it was the last to
    increment (the general path).
Checked under random interleavings: the output is written exactly once, by a piece that saw every partial it used, the
output is the fold of ALL pieces, and the counter is back at zero for the next launch.
"""
import random
import threading
import time

import pytest


class Pair:
    def __init__(self, pieces):
        self.n = pieces
        self.counter = 0
        self.lock = threading.Lock()          # the atomicAdd
        self.partial = [None] * pieces        # published partials (the workspace slots)
        self.outputs = []                     # (writer, value)
        self.errors = []

    def atomic_inc(self):
        with self.lock:
            old = self.counter
            self.counter = old + 1
            return old


def piece(pair, idx, value, rnd, start_delay, length):
    time.sleep(start_delay)
    seen = pair.counter                       # read when the segment begins
    time.sleep(length * 0.8)
    seen = pair.counter                       # ... and again one tile before it ends
    time.sleep(length * 0.2 + rnd.random() * 0.001)
    if seen == pair.n - 1:
        # direct path: every other piece has published (publication precedes the increment)
        others = [pair.partial[j] for j in range(pair.n) if j != idx]
        if any(o is None for o in others):
            pair.errors.append("piece %d merged before a partner had published" % idx)
            return
        pair.outputs.append((idx, sum(others) + value))
        pair.counter = 0
    else:
        pair.partial[idx] = value             # partial, then fence, then the counter
        old = pair.atomic_inc()
        if old == pair.n - 1:
            if any(o is None for o in pair.partial):
                pair.errors.append("last arriver %d found a slot empty" % idx)
                return
            pair.outputs.append((idx, sum(pair.partial)))
            pair.counter = 0


@pytest.mark.parametrize("pieces", [2, 3, 5])
def test_pair_is_merged_exactly_once_from_all_pieces(pieces):
    rnd = random.Random(pieces)
    for trial in range(60):
        pair = Pair(pieces)
        values = [float(1 << i) for i in range(pieces)]      # every subset sums to a different number
        shape = trial % 3
        threads = []
        for i in range(pieces):
            if shape == 0:    # the usual picture: piece 0 ends the kernel, its partners were first segments
                start, length = (0.004, 0.002) if i == 0 else (0.0, 0.001)
            elif shape == 1:  # pairs longer than a CTA's share: everything overlaps
                start, length = rnd.random() * 0.001, 0.002 + rnd.random() * 0.001
            else:             # all at once
                start, length = 0.0, 0.001
            threads.append(threading.Thread(target=piece, args=(pair, i, values[i], random.Random(trial * 10 + i), start, length)))
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        assert pair.errors == []
        assert len(pair.outputs) == 1, pair.outputs
        assert pair.outputs[0][1] == sum(values)
        assert pair.counter == 0


def test_direct_path_is_taken_when_partners_are_long_done():
    pair = Pair(2)
    ts = [threading.Thread(target=piece, args=(pair, 0, 1.0, random.Random(0), 0.01, 0.002)),
          threading.Thread(target=piece, args=(pair, 1, 2.0, random.Random(1), 0.0, 0.001))]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert pair.outputs == [(0, 3.0)] and pair.partial[0] is None and pair.counter == 0
