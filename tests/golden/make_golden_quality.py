"""Generates tests/golden/ref_quality.json by EXECUTING the reference's own QualityScoreGenerationIterator over
Illumina18QualityScoreCodec<float> (src/index/quality_score_iterator.hpp, quality_scores.hpp) through oracle/_ref. Run in
the authoring container only:  make -C oracle && python tests/golden/make_golden_quality.py
Content: the 96-entry decode table (float bit patterns) and, for K in (15, 21, 31, 35, 47, 63), the scores (float bit
patterns) of every K-window of the quality lines of the first records of three FASTQ fixtures and of one seeded random line
that holds every quality byte 33..126."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_util as ou  # noqa: E402

KS = (15, 21, 31, 35, 47, 63)


def quality_lines(buf, limit):
    lines = [ln for ln in bytes(buf).replace(b"\r", b"\n").split(b"\n") if ln]
    return lines[3::4][:limit]


def main():
    ref = ou.ref_lib("ndebug")
    assert ref is not None and hasattr(ref, "kmiref_quality"), "build oracle/_ref first (make -C oracle)"
    lut = np.zeros(96, dtype=np.float32)
    ref.kmiref_quality_lut(ou._ptr(lut))
    rng = np.random.default_rng(2024)
    lines = {"random": [bytes(rng.permutation(np.arange(33, 127, dtype=np.uint8))) + bytes(rng.integers(35, 75, 200, dtype=np.uint8))]}
    for fx in ("natural.fastq", "test.small.fastq", "test.unitiq1.short.fastq"):
        lines[fx] = quality_lines(ou.read_fixture(fx), 3)
    out = {"lut_bits": [int(x) for x in lut.view(np.uint32)], "lines": {k: [ln.decode("latin1") for ln in v] for k, v in lines.items()}, "scores": {}}
    for k in KS:
        per = {}
        for name, lns in lines.items():
            rows = []
            for ln in lns:
                q = np.frombuffer(ln, dtype=np.uint8).copy()
                sc = np.zeros(max(q.size, 1), dtype=np.float32)
                m = ref.kmiref_quality(k, ou._ptr(q), q.size, ou._ptr(sc))
                rows.append([int(x) for x in sc[:m].view(np.uint32)])
            per[name] = rows
        out["scores"][str(k)] = per
    json.dump(out, open(os.path.join(HERE, "ref_quality.json"), "w"))
    print("wrote ref_quality.json")


if __name__ == "__main__":
    main()
