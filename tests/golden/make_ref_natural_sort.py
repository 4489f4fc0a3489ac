"""Writes tests/golden/ref_natural_sort.json.gz: SI::natural::compare of the reference's vendored dependencies/NaturalSort
(compiled as it is into oracle/_ref/libref_json.so by oracle/build_ref.sh) on pairs of frame names - hand-picked edge cases
plus seeded random names over an alphabet that makes digit runs, dots, spaces, case and non-ASCII bytes meet.
Run here (needs /root/reference): python tests/golden/make_ref_natural_sort.py"""
import ctypes
import gzip
import json
import os
import random

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_json.so"))
lib.ref_natural_less.argtypes = [ctypes.c_char_p, ctypes.c_char_p]


def main():
    names = ["img-1", "img1", "b1", "B2", "a1.5", "a10", "a.5", "a.10", "a.50", "a.5000", "x 9", "x  10", "x   10", "été", "view", "Été", "写真_2", "写真_10", "a", "a1",
             "a01", "a001", "frame_2", "frame_10", "frame_010", "a ", "a  b", "a b", "1", "/", "images/0001.jpg", "images/1.jpg", "images/10.jpg", "images\\2.jpg",
             "./train/r_9", "./train/r_10", "./train/r_100", "r_9.5", "r_9.25", "r_9.250", "3.14", "3.2", "3.", ".3", "..3", "v1.2.10", "v1.2.9", "v1.10.1", "Z", "z", "_", "~",
             "0", "00", "0a", "00a", "9" * 25, "1" + "0" * 25, "a9" * 3, ""]
    rng = random.Random(7)
    alphabet = list("aAbBzZ0123459 ._-/") + ["é", "写"]
    for _ in range(400):
        names.append("".join(rng.choice(alphabet) for _ in range(rng.randint(1, 9))))
    pairs = [(a, b) for a in names[:60] for b in names[:60]]
    for _ in range(6000):
        pairs.append((rng.choice(names), rng.choice(names)))
    # the reference walks past the end of both strings when they are exhausted together by the space skipping (undefined
    # behaviour, e.g. "a  " against "a  "): such pairs are left out
    def defined(a, b):
        return not (a.rstrip(" ") != a and b.rstrip(" ") != b)
    out = [[a, b, int(lib.ref_natural_less(a.encode(), b.encode()))] for a, b in pairs if defined(a, b)]
    with gzip.GzipFile(os.path.join(HERE, "ref_natural_sort.json.gz"), "wb", mtime=0) as f:
        f.write(json.dumps({"made_with": "SI::natural::compare<std::string>, dependencies/NaturalSort/natural_sort.hpp", "pairs": out}, ensure_ascii=True, separators=(",", ":")).encode())
    print(len(out), "pairs,", sum(p[2] for p in out), "less")


if __name__ == "__main__":
    main()
