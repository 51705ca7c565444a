#!/usr/bin/env python
"""Thread model of the arch5 Task hand-over protocol (engine.cc: send_task / recv_begin / release_stage / extract_loop):
samplers with a ring of kStagePerSampler stage buffers, a ticketed descriptor ring, one synchronous trainer and one
prefetching trainer with a bounded pool and a slow consumer.  `hold_stage_while_dequeuing=True` is the first version of
the extract loop (a prefetched Task kept its stage until its gather): it deadlocks when the sampler that must produce
the next Task has cycled back to that stage.  The shipped rule - release before every blocking dequeue - never does.

  python tools/arch5_queue_model.py            # prints deadlocks / trials for both variants
"""
import random
import sys
import threading
import time

STAGES, DEPTH = 4, 16
TIMEOUT = 0.5  # a wait this long means nobody will ever post: deadlock (tests raise it)


class Deadlock(Exception):
    pass


def acquire(sem):
    if not sem.acquire(timeout=TIMEOUT):
        raise Deadlock()


def run_trial(seed, hold_stage_while_dequeuing, samplers=2, tasks_per_sampler=8):
    rnd = random.Random(seed)
    total = samplers * tasks_per_sampler
    stage_free = [[threading.Semaphore(1) for _ in range(STAGES)] for _ in range(samplers)]
    filled = [threading.Semaphore(0) for _ in range(DEPTH)]
    vacant = [threading.Semaphore(1) for _ in range(DEPTH)]
    desc = [None] * DEPTH
    lock = threading.Lock()
    cnt = {"send": 0, "recv": 0}
    failed, delivered = [], []
    # log-uniform delays, 10 us .. 10 ms: the bad interleaving needs a slow consumer behind the prefetching trainer while the
    # other trainer drains the queue and one sampler has already finished
    d_send = [10 ** rnd.uniform(-5, -2) for _ in range(samplers)]
    d_consume = [10 ** rnd.uniform(-5, -2) for _ in range(2)]

    def ticket(which):
        with lock:
            t = cnt[which]
            cnt[which] += 1
            return t

    def guard(fn):
        def w(*a):
            try:
                fn(*a)
            except Deadlock:
                failed.append(fn.__name__)
        return w

    @guard
    def sampler(s):
        for i in range(tasks_per_sampler):
            k = i % STAGES
            acquire(stage_free[s][k])
            time.sleep(d_send[s] * rnd.random())
            slot = ticket("send") % DEPTH
            acquire(vacant[slot])
            desc[slot] = (s, k, i)
            filled[slot].release()

    def begin():
        slot = ticket("recv") % DEPTH
        acquire(filled[slot])
        d = desc[slot]
        vacant[slot].release()
        return {"d": d, "held": True}

    def release(p):
        if p["held"]:
            p["held"] = False
            stage_free[p["d"][0]][p["d"][1]].release()

    @guard
    def trainer_sync(mine):
        for _ in range(mine):
            p = begin()
            release(p)
            delivered.append(p["d"])
            time.sleep(d_consume[0] * rnd.random())

    @guard
    def trainer_prefetch(mine):
        pool, cv = [], threading.Condition()

        def consumer():
            for _ in range(mine):
                with cv:
                    if not cv.wait_for(lambda: pool, timeout=TIMEOUT * 4):
                        return
                    delivered.append(pool.pop(0))
                    cv.notify_all()
                time.sleep(d_consume[1] * rnd.random())

        c = threading.Thread(target=consumer)
        c.start()
        nxt = None
        for i in range(mine):
            if nxt is not None and not hold_stage_while_dequeuing:
                release(nxt)
            with cv:
                if not cv.wait_for(lambda: len(pool) < 2, timeout=TIMEOUT * 4):
                    raise Deadlock()
            cur = nxt if nxt is not None else begin()
            nxt = None
            if not hold_stage_while_dequeuing:
                release(cur)
            if i + 1 < mine:
                nxt = begin()
            release(cur)
            with cv:
                pool.append(cur["d"])
                cv.notify_all()
        c.join()

    th = [threading.Thread(target=sampler, args=(s,)) for s in range(samplers)]
    th += [threading.Thread(target=trainer_sync, args=(total // 2,)), threading.Thread(target=trainer_prefetch, args=(total - total // 2,))]
    for t in th:
        t.start()
    for t in th:
        t.join()
    ok = not failed and sorted(delivered) == sorted((s, i % STAGES, i) for s in range(samplers) for i in range(tasks_per_sampler))
    return ok


def main():
    trials = int(sys.argv[1]) if len(sys.argv) > 1 else 60
    for hold in (True, False):
        bad = sum(not run_trial(seed, hold) for seed in range(trials))
        print(f"hold_stage_while_dequeuing={hold}: {bad} deadlocked / {trials} trials")


if __name__ == "__main__":
    main()
