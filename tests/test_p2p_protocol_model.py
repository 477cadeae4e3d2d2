"""Model of the NVLink peer-window exchange protocol (cuda_samples_b200/csrc/p2p.cuh) -- CPU only.

The windows have two slots per source rank (parity of the exchange number).  Rank r publishes word {s, data} of exchange
s into slot [s & 1][r] of EVERY window and polls its OWN window for the words it needs.  Slot s & 1 is rewritten by
exchange s + 2, so the protocol is only safe if no rank can start exchange s + 2 while a peer still has to read words of
exchange s.  p2p.cuh gets that from one rule: in every exchange every rank waits for every rank's word.

The model below runs the ranks as independent step machines under random schedules and checks both directions of that
rule: "read everybody" never loses a word; "read only the ranks below" (what a scan's carry needs -- the first version
of scan_tile_sums_kernel did exactly that) lets rank 0 run ahead and overwrite words its peers have not read yet.  On
the GPU a lost word is a poll that never sees its exchange number: a timeout and a sticky error count, never a wrong sum.
"""
import random

import pytest


def run(world, exchanges, readers, schedule_seed, bias_rank=None):
    """-> number of lost words.  readers(r) = the source ranks rank r waits for in every exchange."""
    rng = random.Random(schedule_seed)
    # window[r][parity][src] = exchange number of the word rank src last stored there (0: nothing yet)
    window = [[[0] * world for _ in range(2)] for _ in range(world)]
    state = [{"s": 1, "published": False, "todo": None} for _ in range(world)]
    lost = 0
    live = set(range(world))
    while live:
        ranks = sorted(live)
        weights = [8 if r == bias_rank else 1 for r in ranks]  # a rank the scheduler favours (it "runs ahead")
        r = rng.choices(ranks, weights)[0]
        st = state[r]
        s = st["s"]
        if not st["published"]:
            for peer in range(world):
                window[peer][s & 1][r] = s
            st["published"] = True
            st["todo"] = list(readers(r))
            continue
        if st["todo"]:
            src = st["todo"][0]
            seen = window[r][s & 1][src]
            if seen == s:
                st["todo"].pop(0)
            elif seen > s:  # overwritten by a later exchange before it was read: the poll can never succeed
                lost += 1
                st["todo"].pop(0)
            continue  # (seen < s: not there yet, poll again later)
        st["s"], st["published"] = s + 1, False
        if st["s"] > exchanges:
            live.discard(r)
    return lost


@pytest.mark.parametrize("world", [2, 4, 8])
def test_every_rank_waits_for_every_rank_never_loses_a_word(world):
    for seed in range(40):
        assert run(world, 12, lambda r: range(world), seed, bias_rank=seed % world) == 0


@pytest.mark.parametrize("world", [2, 4, 8])
def test_waiting_for_the_lower_ranks_only_lets_rank_0_lap_its_peers(world):
    """The scan's carry needs the lower ranks' totals only -- but an exchange that reads just those is not a barrier."""
    losses = sum(run(world, 12, lambda r: range(r), seed, bias_rank=0) for seed in range(40))
    assert losses > 0
