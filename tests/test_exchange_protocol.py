"""The exchange protocol of pmtech_b200/csrc/pm_exchange.cu as a model: every rank is an IN-ORDER queue of operations (on the
boxes this was developed on a waiting stream holds back everything enqueued behind it on its device), the operations are
the ones pmb_exchange_launch / pmb_exchange_finish enqueue, in their order, and the ranks advance in random interleavings.

    launch(e)   pack(e) publish(e) | waitREADY(e-1) push(e-1) waitDONE(e-1)      (the previous frame's phases 2-3)
    finish      waitREADY(e) push(e) waitDONE(e)

Checked for every interleaving tried: nobody waits for ever; push(e) on every rank reads, from every peer, the table
slot that peer published for epoch e (two buffer sets, reused every other frame).  The model also holds the protocol
as it was before the DONE barrier was extended to every rank — lists gathered on rank 0, only rank 0 waiting for DONE —
and shows the failure seen on 8 GPUs: a fast rank publishes epoch e + 2 over a slot a slower rank still needs."""
import random

import pytest


def simulate(world, frames, seed, done_wait_on, slow=None, max_steps=200000):
    """done_wait_on: set of ranks whose phase 3 waits for DONE (the shipped protocol: all of them).
    Returns (ok, why).  slow: rank -> probability of being skipped when picked (a rank that lags)."""
    rng = random.Random(seed)
    table = [[[0] * world for _ in range(2)] for _ in range(world)]   # table[at][set][from] = epoch of the published slot
    ready = [[[0] * world for _ in range(2)] for _ in range(world)]
    done = [[[0] * world for _ in range(2)] for _ in range(world)]
    queues = []
    for r in range(world):
        q = []
        for e in range(1, frames + 1):
            q.append(("frame", e))
            q.append(("publish", e))
            if e > 1:
                q += [("wait_ready", e - 1), ("push", e - 1)] + ([("wait_done", e - 1)] if r in done_wait_on else [])
        q += [("wait_ready", frames), ("push", frames)] + ([("wait_done", frames)] if r in done_wait_on else [])
        queues.append(q)
    pc = [0] * world
    for _ in range(max_steps):
        if all(pc[r] == len(queues[r]) for r in range(world)):
            return True, "done"
        runnable = []
        for r in range(world):
            if pc[r] == len(queues[r]):
                continue
            op, e = queues[r][pc[r]]
            b = e & 1
            if op == "wait_ready" and any(ready[r][b][q] < e for q in range(world)):
                continue
            if op == "wait_done" and any(done[r][b][q] < e for q in range(world)):
                continue
            runnable.append(r)
        if not runnable:
            return False, "deadlock at " + str([queues[r][pc[r]] if pc[r] < len(queues[r]) else None for r in range(world)])
        r = rng.choice(runnable)
        if slow and rng.random() < slow.get(r, 0.0):
            continue
        op, e = queues[r][pc[r]]
        b = e & 1
        if op == "publish":
            for q in range(world):
                table[q][b][r] = e
                ready[q][b][r] = e
        elif op == "push":
            stale = [q for q in range(world) if table[r][b][q] != e]
            if stale:
                return False, f"rank {r} push({e}) reads the slots of ranks {stale} published for epoch {[table[r][b][q] for q in stale]}"
            for q in range(world):
                done[q][b][r] = e
        pc[r] += 1
    return False, "step budget exhausted"


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_the_shipped_protocol_never_blocks_and_never_reads_a_stale_table(world):
    for seed in range(60):
        slow = {0: 0.9} if seed % 3 == 0 else ({world - 1: 0.8} if seed % 3 == 1 else None)   # the receiving rank lags; or the last one
        ok, why = simulate(world, frames=7, seed=seed, done_wait_on=set(range(world)), slow=slow)
        assert ok, (world, seed, why)


def test_waiting_for_done_on_the_receiver_alone_lets_a_fast_rank_overwrite_the_table():
    """the protocol as first written (only rank 0, which receives the lists, waits for DONE): with a lagging rank 0 some
    interleaving makes a peer publish epoch e + 2 into the slot rank 0 still needs for epoch e"""
    bad = 0
    for seed in range(60):
        ok, why = simulate(8, frames=7, seed=seed, done_wait_on={0}, slow={0: 0.9})
        bad += 0 if ok else 1
    assert bad > 0
