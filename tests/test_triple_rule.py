"""CPU suite: the combinatorial rule the bitset kernel relies on for paths that visit a node more
than once (gfalign_b200/csrc/kernels.cu build_tables_kernel):

    B occurs contiguously in A   <=>   every consecutive pair of B is taken by A
                                       and no consecutive triple of B is a *broken triple* of A

where the broken triples of A are (pred of one visit of x, x, succ of another visit of x) when that
triple never occurs contiguously, plus the real triples around every pair A takes more than once.
Checked exhaustively-at-random on small alphabets (orientation is just a larger alphabet)."""
import random


def _occurs(A, B):
    m = len(B)
    return any(A[i:i + m] == B for i in range(len(A) - m + 1))


def _broken(A):
    n = len(A)
    real = set(zip(A[:-2], A[1:-1], A[2:]))
    bt = set()
    for i in range(n):
        for v2 in range(n):
            if v2 == i or A[v2] != A[i]:
                continue
            if i >= 1 and v2 + 1 < n and (A[i - 1], A[i], A[v2 + 1]) not in real:
                bt.add((A[i - 1], A[i], A[v2 + 1]))
            if i + 1 < n and v2 + 1 < n and A[v2 + 1] == A[i + 1]:
                if i >= 1:
                    bt.add((A[i - 1], A[i], A[i + 1]))
                if i + 2 < n:
                    bt.add((A[i], A[i + 1], A[i + 2]))
    return set(zip(A[:-1], A[1:])), bt


def test_pairs_and_no_broken_triple_imply_occurrence():
    rnd = random.Random(1)
    certified = 0
    for _ in range(60000):
        alpha = rnd.choice([2, 3, 4, 5])
        A = [rnd.randrange(alpha) for _ in range(rnd.randint(2, 14))]
        B = [rnd.randrange(alpha) for _ in range(rnd.randint(2, 7))]
        if len(B) > len(A):
            continue
        pairs, bt = _broken(A)
        cand = all(p in pairs for p in zip(B[:-1], B[1:]))
        if _occurs(A, B):
            assert cand                      # necessary
        if cand and not any(t in bt for t in zip(B[:-2], B[1:-1], B[2:])):
            assert _occurs(A, B), (A, B)     # sufficient
            certified += 1
    assert certified > 1000
