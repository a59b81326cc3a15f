"""The division-free work-item walk of the dual-query-tile prefill kernel (csrc/prefill_attn_dq.cu:
ItemIter, it_begin / it_next) restated in Python and checked against the closed-form item decode it
replaced (k -> chunk of requests, query tile descending, request, pair of q heads).  The kernel itself
is covered by the GPU parity tests; this pins the index arithmetic on the CPU, including chunks whose
last one is short, fewer items than CTAs, and grid sizes that are not multiples of the pair count."""
import random


def closed_form(k, q_tiles, pairs, chunk_reqs, num_reqs):
    per_chunk = q_tiles * pairs * chunk_reqs
    c, k2 = divmod(k, per_chunk)
    r0 = c * chunk_reqs
    pq = pairs * min(chunk_reqs, num_reqs - r0)
    lv, rem = divmod(k2, pq)
    return q_tiles - 1 - lv, r0 + rem // pairs, rem % pairs


def walk(block, grid, q_tiles, pairs, chunk_reqs, num_reqs):
    """it_begin() for CTA `block`, then it_next() until the end: yields (k, (qt, request, pair))."""
    total = q_tiles * pairs * num_reqs
    step_rr = grid // pairs
    step_pr = grid - step_rr * pairs
    k = block
    if k >= total:
        return
    per_chunk = q_tiles * pairs * chunk_reqs
    c, k2 = divmod(k, per_chunk)
    r0 = c * chunk_reqs
    nreq = min(chunk_reqs, num_reqs - r0)
    lv, rem = divmod(k2, pairs * nreq)
    qt = q_tiles - 1 - lv
    rr, pr = divmod(rem, pairs)
    while True:
        yield k, (qt, r0 + rr, pr)
        k += grid
        if k >= total:
            return
        pr += step_pr
        rr += step_rr
        if pr >= pairs:
            pr -= pairs
            rr += 1
        while rr >= nreq:
            rr -= nreq
            qt -= 1
            if qt < 0:
                qt = q_tiles - 1
                r0 += nreq
                nreq = min(chunk_reqs, num_reqs - r0)


def test_item_iterator_matches_closed_form():
    rng = random.Random(1)
    steps = 0
    for _ in range(400):
        q_tiles = rng.randint(1, 20)
        pairs = rng.choice([1, 2, 3, 4, 6, 8, 16, 20])
        num_reqs = rng.randint(1, 70)
        chunk_reqs = rng.randint(1, num_reqs)
        grid = rng.choice([1, 3, 16, 37, 147, 148])
        for block in rng.sample(range(grid), min(grid, 3)):
            for k, got in walk(block, grid, q_tiles, pairs, chunk_reqs, num_reqs):
                assert got == closed_form(k, q_tiles, pairs, chunk_reqs, num_reqs), \
                    (k, q_tiles, pairs, chunk_reqs, num_reqs, grid)
                steps += 1
    assert steps > 100000


def test_item_iterator_bench_shape_covers_every_item_once():
    # 64 prompts x 1024 tokens, Hq 32: 8 query tiles x 16 pairs x 64 requests over 148 CTAs
    q_tiles, pairs, num_reqs, chunk_reqs, grid = 8, 16, 64, 64, 148
    seen = set()
    for block in range(grid):
        for k, pos in walk(block, grid, q_tiles, pairs, chunk_reqs, num_reqs):
            assert pos not in seen
            seen.add(pos)
    assert len(seen) == q_tiles * pairs * num_reqs
