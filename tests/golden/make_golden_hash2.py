"""Generates tests/golden/ref_kat_hash2.json by EXECUTING the reference's own identity / cpp_std hash classes
(src/index/kmer_hash.hpp:157-232) through oracle/_ref (built by oracle/Makefile from /root/reference), constructed as
KeyToRank constructs them: prefix bits = ceilLog2(comm_size). Run in the authoring container only:
    make -C oracle && python tests/golden/make_golden_hash2.py
Keys: the canonical k-mers of tests/golden/ref_kat.json (same order). Per k: store hashes, and dist hashes for
p in (2, 3, 8, 64, 200) ranks (p = 1 leaves cpp_std's shift undefined for K > 32 and maps everything to rank 0 anyway).
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_util as ou  # noqa: E402

PS = (2, 3, 8, 64, 200)


def main():
    ref = ou.ref_lib("ndebug")
    assert ref is not None and hasattr(ref, "kmiref_hash_prefix"), "build oracle/_ref first (make -C oracle)"
    kat = json.load(open(os.path.join(HERE, "ref_kat.json")))
    out = {"ranks": list(PS), "k": {}}
    for ks, e in kat["k"].items():
        k = int(ks)
        can = np.array([int(x, 16) for x in e["canonical"]], dtype=np.uint64)
        n = can.size // ou.nwords(k)
        ent = {}
        for name, kind in (("identity", 0), ("std", 2)):
            h = np.zeros(n, dtype=np.uint64)
            ref.kmiref_hash_prefix(k, kind, 0, ou._ptr(can), ou._ptr(h), n)
            ent[name + "_store"] = ["%016x" % int(x) for x in h]
            for p in PS:
                ref.kmiref_hash_prefix(k, kind + 1, ou.ceil_log2(p), ou._ptr(can), ou._ptr(h), n)
                ent["%s_dist_p%d" % (name, p)] = ["%016x" % int(x) for x in h]
        out["k"][ks] = ent
    with open(os.path.join(HERE, "ref_kat_hash2.json"), "w") as f:
        json.dump(out, f, indent=0)
    print("wrote ref_kat_hash2.json")


if __name__ == "__main__":
    main()
