"""Shared helpers for the test-suite."""
import base64
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLDEN = os.path.join(HERE, "golden")
FIXTURES = os.path.join(GOLDEN, "fixtures")


def golden_cases():
    with open(os.path.join(GOLDEN, "cases.json")) as fh:
        cases = json.load(fh)
    for c in cases:
        c["xgfa"] = base64.b64decode(c["xgfa_b64"]) if c.get("xgfa_b64") else None
    return cases


def case_id(c):
    return c["fixture"].replace(".fasta", "") + ":" + " ".join(c["flags"])


def elastic_cases():
    return [c for c in golden_cases() if "--elastic" in c["flags"] and "--gfa" in c["flags"]]


def nonelastic_cases():
    return [c for c in golden_cases() if c["flags"][0].startswith("--gap-limit")]


def flags_to_opts(flags):
    return dict(ignore=b"N" if "--ignore-chars=N" in flags else b"", tricks_off="--disable-elastic-tricks" in flags,
                paths="-p" in flags)


def stat_lines(c):
    out = {}
    for l in c["stderr"]:
        for k in ("Optimal score: ", "Number of segments: ", "#nodes=", "total length of node labels=", "#founders=", "#edges="):
            if l.startswith(k) and k not in out:  # make_efg repeats two of them with 0 before it crashes (fbg.cpp:1001-1002)
                out[k] = int(l[len(k):])
        if l.startswith("MSA contains") and "gaps" in l:
            out["gaps"] = int(l.split()[2])
        if l.startswith("MSA contains") and "ignore" in l:
            out["toignore"] = int(l.split()[2])
        if "optimal segment length" in l:
            out["L"] = int(l.split("= ")[1].split(")")[0])
    return out
