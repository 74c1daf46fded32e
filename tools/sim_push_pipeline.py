#!/usr/bin/env python
"""Queueing model behind DESIGN.md 6.a: how much of the owners' time does the push transport lose to the uneven number of
pages per owner and round, as a function of how many multi-gets every requester keeps in flight?

n GPUs; every round each requester draws 256 uniform owners (its own pages cost a fraction of a remote one).  An owner
serves round k as one unit: it starts when it has finished round k-1 AND every requester's posting of round k is in,
and takes (pages asked of it) time units.  A requester posts round k when its round k-depth has completed, i.e. when
every owner has finished that round.  Efficiency = ideal time (mean load) / time of the slowest requester.

  python tools/sim_push_pipeline.py            # prints a small table (no GPU needed)

Result (n = 8): one multi-get in flight 0.91, two 0.99, more no better -- the loss of a lock-step protocol is the +9 % of
the most loaded owner paid EVERY round; with two in flight the owners drift apart and the spread averages out.  (The
measured gain from more than two in flight, DESIGN.md 5, comes from latencies this model leaves out: the post -> serve ->
done -> wait hand-overs of a round.)"""
import numpy as np


def efficiency(n=8, rounds=400, depth=2, seed=0, local_cost=0.15, per_requester=False):
    rng = np.random.default_rng(seed)
    cnt = rng.multinomial(256, [1.0 / n] * n, size=(rounds, n)).astype(float)  # [round][requester][owner]
    for k in range(rounds):
        for r in range(n):
            cnt[k, r, r] *= local_cost
    done = np.zeros((rounds, n))
    owner_free = np.zeros(n)
    for k in range(rounds):
        post = np.zeros(n) if k < depth else done[k - depth].copy()
        fin = np.zeros((n, n))
        for o in range(n):
            if per_requester:  # an owner that serves requesters one by one, as their postings arrive
                t = owner_free[o]
                for r in np.argsort(post):
                    t = max(t, post[r]) + cnt[k, r, o]
                    fin[r, o] = t
            else:              # the serve kernel: waits for every posting of the round, then serves them all
                t = max(owner_free[o], post.max()) + cnt[k, :, o].sum()
                fin[:, o] = t
            owner_free[o] = t
        done[k] = np.maximum(fin.max(axis=1), post)
    ideal = rounds * (256.0 * (n - 1) / n + local_cost * 256.0 / n)
    return ideal / done[-1].max()


if __name__ == "__main__":
    print("GPUs  in flight  serve kernel  per-requester service")
    for n in (2, 4, 8):
        for depth in (1, 2, 3, 4, 6):
            a = np.mean([efficiency(n=n, depth=depth, seed=s) for s in range(3)])
            b = np.mean([efficiency(n=n, depth=depth, seed=s, per_requester=True) for s in range(3)])
            print("%4d  %9d  %12.3f  %21.3f" % (n, depth, a, b))
