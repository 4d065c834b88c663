"""Golden vectors of the REFERENCE's FASTQParser::find_first_record (src/io/fastq_loader.hpp:269-357), executed here through
oracle/_ref/ref_t1 (mode first_record: the reference's parser class compiled from /root/reference over the single-rank mpi.h
of oracle/mpishim). Per FASTQ fixture: (search begin, search end, result) for a grid of cuts x window sizes — the whole rest
of the file, and windows of 64 / 300 / 1000 / 5000 bytes, where the reference throws when the window holds four line starts
that are not FASTQ markers (result 2^64 - 1) — and the block boundaries of p = 2, 3, 8, 64 ranks as
partitioned_file<posix_file, FASTQParser>::read_file finds them (file.hpp:1342-1431). Committed as
tests/golden/ref_first_record.npz; tests/test_partition.py holds kmi_fastq_find_first_record, the C oracle's restatement and
kmi_file_block to every row. Run in the authoring container (needs /root/reference): python tests/golden/make_golden_first_record.py"""
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
EXE = os.path.join(ROOT, "oracle", "_ref", "ref_t1")
FASTQ = ["natural.fastq", "natural.withN.fastq", "test.debruijn.small.fastq", "test.debruijn.tiny.fastq", "test.medium.fastq",
         "test.small.fastq", "test.unitiq1.fastq", "test.unitiq1.short2.fastq", "test.unitiq1.short.fastq", "test.unitiq2.fastq",
         "test.unitiqs.fastq"]


def synthetic():
    """texts that are not among the fixtures (tests/test_partition.py rebuilds the same bytes): quality lines that begin with
    '@' and '+', the ambiguity the four-line vote exists for; and lines without any FASTQ marker, where windows that do not
    reach the end of the file make the reference throw"""
    rec = []
    for i in range(60):
        seq = "ACGTTGCAAGGCTTAACCGG"[i % 7:] + "ACGT" * (3 + i % 5)
        qual = ("@+"[i % 2]) + "I" * (len(seq) - 1)
        rec.append("@r%d\n%s\n+\n%s\n" % (i, seq, qual))
    junk = "".join("line %d of no format at all\n" % i for i in range(120))
    return {"synthetic.at_quality.fastq": "".join(rec).encode(), "synthetic.junk.fastq": junk.encode()}


def main():
    if not os.path.exists(EXE):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])
    out = {}
    with tempfile.TemporaryDirectory() as tmp:
        paths = {fx: os.path.join(HERE, "data", fx) for fx in FASTQ}
        for name, text in synthetic().items():
            paths[name] = os.path.join(tmp, name)
            open(paths[name], "wb").write(text)
        for fx in paths:
            dump = os.path.join(tmp, "fr.bin")
            subprocess.check_call([EXE, "first_record", "0", paths[fx], dump], stdout=subprocess.DEVNULL)
            a = np.fromfile(dump, dtype=np.uint64)
            n, nw = int(a[0]), int(a[1])
            assert nw == 2 and a.size == 2 + 3 * n
            out[fx] = a[2:].reshape(n, 3)
            thrown = int((out[fx][:, 2] == np.uint64(2**64 - 1)).sum())
            print("%-28s %5d calls, %4d where the reference throws" % (fx, n, thrown))
    np.savez_compressed(os.path.join(HERE, "ref_first_record.npz"), **out)


if __name__ == "__main__":
    sys.exit(main())
