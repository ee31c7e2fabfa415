#!/usr/bin/env python
"""Copies the reference's 42 nnet_conf/*.json model configs (egs/*/v*/nnet_conf) into tests/golden/nnet_conf/ as parser
fixtures: SURVEY.md section 8(b) requires that every one of them loads unchanged through Params and through
xv_create's own JSON reader (comment-like "Note" keys, ints, floats in exponent form, booleans, lists).

  python tests/golden/make_nnet_conf_fixtures.py        (needs /root/reference; the copies are committed)

File name = <recipe>_<version>__<original name>.  These are configuration DATA (the schema the drop-in must accept),
not code."""
import glob
import os
import shutil

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    dst = os.path.join(HERE, "nnet_conf")
    os.makedirs(dst, exist_ok=True)
    n = 0
    for path in sorted(glob.glob(os.path.join(REF, "egs", "*", "*", "nnet_conf", "*.json"))):
        parts = path.split(os.sep)
        name = "%s_%s__%s" % (parts[-4], parts[-3], parts[-1])
        shutil.copyfile(path, os.path.join(dst, name))
        n += 1
    print("copied %d configs" % n)


if __name__ == "__main__":
    main()
