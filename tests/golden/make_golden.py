"""Generates tests/golden/loader_golden.json by running the REFERENCE's own Matrix-Market loader
(/root/reference/ReadMatrixMarket, compiled into oracle/_ref by oracle/Makefile) on the three
fixtures the reference ships (copied here as data: one.mtx, b1_ss.mtx, bcspwr01.mtx) and on a set
of small hand-made files that exercise its quirks (duplicates + unsorted rows -> its quicksort's
order, 0-based files, pattern / integer / complex fields, symmetric expansion defect, rectangular,
transpose).  Run in the build container only (the reference tree is not on the GPU box):

    make -C oracle && python tests/golden/make_golden.py
"""
import json
import os
import random
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import orc  # noqa: E402


def write(name, text):
    with open(os.path.join(HERE, name), "w") as f:
        f.write(text)


def make_files():
    rnd = random.Random(20261104)
    # duplicates and unsorted columns, one long row (exercises the reference quicksort's order)
    lines = []
    for i in range(1, 13):
        cols = [rnd.randint(1, 12) for _ in range(rnd.randint(1, 9))]
        if i == 5:
            cols = [rnd.randint(1, 12) for _ in range(40)]
        for c in cols:
            lines.append((i, c, rnd.uniform(-3, 3)))
    rnd.shuffle(lines)
    write("dup_unsorted.mtx", "%%%%MatrixMarket matrix coordinate real general\n%% duplicates, unsorted\n12 12 %d\n" % len(lines)
          + "".join("%d %d %.17g\n" % l for l in lines))
    write("zero_based.mtx", "%%MatrixMarket matrix coordinate real general\n4 4 6\n0 0 1.5\n1 1 -2\n2 2 3.25\n3 3 4\n3 0 0.5\n0 3 1e-3\n")
    write("pattern_general.mtx", "%%MatrixMarket matrix coordinate pattern general\n5 5 7\n1 1\n2 2\n3 3\n4 4\n5 5\n5 1\n2 4\n")
    write("integer_general.mtx", "%%MatrixMarket matrix coordinate integer general\n3 3 4\n1 1 7\n2 2 -3\n3 3 12\n3 1 5\n")
    write("sym_real.mtx", "%%MatrixMarket matrix coordinate real symmetric\n5 5 8\n1 1 4\n2 1 -1\n2 2 4\n3 2 -1.5\n3 3 4\n4 4 4\n5 4 -0.25\n5 5 4\n")
    write("rect.mtx", "%%MatrixMarket matrix coordinate real general\n% rectangular\n3 6 7\n1 6 1\n1 2 2\n2 3 3\n2 1 4\n3 5 5\n3 4 6\n3 6 7\n")
    write("complex_general.mtx", "%%MatrixMarket matrix coordinate complex general\n2 2 3\n1 1 1.5 2.5\n2 2 -1 0\n2 1 3 -4\n")
    write("skew.mtx", "%%MatrixMarket matrix coordinate real skew-symmetric\n3 3 2\n2 1 1.5\n3 2 -2\n")
    write("empty_rows.mtx", "%%MatrixMarket matrix coordinate real general\n6 6 3\n2 2 1\n2 5 2\n6 1 3\n")


CASES = ["one.mtx", "b1_ss.mtx", "bcspwr01.mtx", "dup_unsorted.mtx", "zero_based.mtx", "pattern_general.mtx",
         "integer_general.mtx", "sym_real.mtx", "rect.mtx", "complex_general.mtx", "skew.mtx", "empty_rows.mtx"]


def main():
    assert orc.have_ref(), "oracle/_ref/libloadmatrix_ref.so missing: make -C oracle"
    make_files()
    out = {}
    for name in CASES:
        for base in (0, 1):
            for tr in (0, 1):
                r = orc.ref_load(os.path.join(HERE, name), base, tr)
                key = f"{name}|base={base}|transpose={tr}"
                if r["ok"] != 1:
                    out[key] = {"ok": r["ok"]}
                    continue
                out[key] = {"ok": 1, "m": r["m"], "n": r["n"], "nnz": r["nnz"], "isSymmetric": r["isSymmetric"],
                            "rowptr": r["rowptr"].tolist(), "colidx": r["colidx"].tolist(),
                            "values_hex": [float(v).hex() for v in r["values"]]}
    with open(os.path.join(HERE, "loader_golden.json"), "w") as f:
        json.dump(out, f, indent=0, sort_keys=True)
    print("wrote", len(out), "cases")


if __name__ == "__main__":
    main()
