"""Generates tests/golden/ref_kat.json by EXECUTING the reference's own code (oracle/_ref, built by
oracle/Makefile from /root/reference). Run in the authoring container only:
    make -C oracle && python tests/golden/make_golden.py
The JSON travels to the GPU box; /root/reference does not.

Content, per k in (15, 21, 31, 35, 47, 63): the forward k-mers of a fixed 96-base string plus 64
seeded random k-mers, with the reference's reverse_complement, lex_less, murmur (store/dist), farm
(store/dist) in both build flavours (NDEBUG and DebugTweak), the dense-hash sentinels, and the
sorted (canonical k-mer, count) dump that the reference's Kmer + lex_less + google::dense_hash_map
produce for three fixtures, with the per-rank sizes at p = 1, 2, 3 (KeyToRank = dist_hash % p).
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_util as ou  # noqa: E402

SEQ = b"ACGTACGTTTGACCAGTACGATCGATCGGGATCCAAATTTGGGCCCAAACGTACGTAGCTAGCTAGCATCGNNACGTTGCAAGGCTTAACCGGTTAC"
KS = (15, 21, 31, 35, 47, 63)


def hx(a):
    return ["%016x" % int(x) for x in np.asarray(a, dtype=np.uint64).reshape(-1)]


def fastq_sequences(buf):
    """sequence lines of a well-formed FASTQ fixture (line 2 of every 4 non-empty lines)."""
    lines = [ln for ln in bytes(buf).replace(b"\r", b"\n").split(b"\n") if ln]
    return lines[1::4]


def main():
    ref = ou.ref_lib("ndebug")
    dbg = ou.ref_lib("debug")
    assert ref is not None and dbg is not None, "build oracle/_ref first (make -C oracle)"
    P = ou._ptr
    out = {"seq": SEQ.decode(), "k": {}}
    seq = np.frombuffer(SEQ, dtype=np.uint8).copy()
    for k in KS:
        nw = ou.nwords(k)
        win = np.zeros(len(SEQ) * nw, dtype=np.uint64)
        m = ref.kmiref_windows(k, P(seq), len(SEQ), P(win))
        win = win[:m * nw]
        rnd = np.ascontiguousarray(ou.random_keys(k, 64, seed=k)).reshape(-1)
        keys = np.concatenate([win, rnd])
        n = keys.size // nw
        rc = np.zeros_like(keys); ref.kmiref_revcomp(k, P(keys), P(rc), n)
        can = np.zeros_like(keys); ref.kmiref_canonical(k, P(keys), P(can), n)
        hs = {}
        for name, lib, which in (("murmur_store", ref, 0), ("murmur_dist", ref, 1), ("farm_store", ref, 2),
                                 ("farm_dist", ref, 3), ("farm_debug_store", dbg, 2), ("farm_debug_dist", dbg, 3)):
            h = np.zeros(n, dtype=np.uint64)
            lib.kmiref_hash(k, which, P(can), P(h), n)
            hs[name] = hx(h)
        sent = np.zeros(2 * nw, dtype=np.uint64)
        ref.kmiref_sentinels(k, P(sent))
        entry = {"n_windows": int(m), "keys": hx(keys), "revcomp": hx(rc), "canonical": hx(can), "hash_of_canonical": hs,
                 "sentinels": hx(sent), "count": {}}
        for fx in ("test.small.fastq", "natural.fastq", "natural.withN.fastq"):
            seqs = fastq_sequences(ou.read_fixture(fx))
            cat = np.frombuffer(b"".join(seqs), dtype=np.uint8).copy()
            off = np.zeros(len(seqs) + 1, dtype=np.uint64)
            off[1:] = np.cumsum([len(s) for s in seqs])
            cap = int(cat.size) + 1
            per = {}
            for hk, hname in ((0, "murmur"), (1, "farm")):
                for p in (1, 2, 3):
                    ok = np.zeros(cap * nw, dtype=np.uint64); oc = np.zeros(cap, dtype=np.uint32)
                    sizes = np.zeros(p, dtype=np.uint64)
                    d = ref.kmiref_count_build(k, P(cat), P(off), len(seqs), p, hk, P(ok), P(oc), cap, P(sizes))
                    per["%s_p%d" % (hname, p)] = [int(x) for x in sizes]
                    if hk == 0 and p == 1:
                        kk = ou.shape_keys(ok[:d * nw], k)
                        order = np.lexsort((kk[:, 0], kk[:, 1])) if nw == 2 else np.argsort(kk, kind="stable")
                        entry["count"][fx] = {"distinct": int(d), "keys": hx(kk[order]), "counts": [int(x) for x in oc[:d][order]]}
                    else:
                        assert d == entry["count"][fx]["distinct"]
            entry["count"][fx]["rank_sizes"] = per
        out["k"][str(k)] = entry
    with open(os.path.join(HERE, "ref_kat.json"), "w") as f:
        json.dump(out, f, indent=0, separators=(",", ":"))
    print("wrote ref_kat.json", os.path.getsize(os.path.join(HERE, "ref_kat.json")), "bytes")


if __name__ == "__main__":
    main()
