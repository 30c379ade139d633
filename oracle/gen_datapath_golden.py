#!/usr/bin/env python
"""Writes tests/golden/datapath_vectors.json: small known answers of the data-path math (pattern
words, per-chunk (X,S) checksums) computed by the CPU oracle (oracle/vgpu_oracle.c, cross-checked
against the numpy restatement).  The reference has no such code, so these pin OUR definition
 against accidental change: the oracle and the kernels are both tested against them."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import cpu  # noqa: E402

CB = 2097152
doc = {"chunk_bytes": CB, "patterns": [], "checksums": []}
for pat, npass, seed in ((0, 0, 0), (0, 5, 0), (0, 63, 0), (1, 0, 0), (1, 7, 0), (1, 42, 0xABCDEF)):
    w = cpu.fill(8, pat, npass, seed)
    far = cpu.fill(4, pat, npass, seed, first=(40 << 30) // 8 - 4)  # last words of a 40 GiB buffer
    doc["patterns"].append({"pattern": pat, "pass": npass, "seed": seed, "first_words": [hex(int(x)) for x in w],
                            "words_before_40GiB": [hex(int(x)) for x in far]})
for npass, nchunks, chunk_bytes in ((42, 4, CB), (3, 3, 4096), (9, 2, CB + 4096), (1, 5, 32)):
    words = cpu.fill(nchunks * chunk_bytes // 8, cpu.PAT_RANDOM, npass, 0)
    n = chunk_bytes // 8
    xs = []
    for c in range(nchunks):
        x, s = cpu.checksum(words[c * n:(c + 1) * n])
        assert (x, s) == cpu.checksum_numpy(words[c * n:(c + 1) * n])
        xs.append([hex(x), hex(s)])
    doc["checksums"].append({"source": f"random pattern, pass {npass}, seed 0", "pass": npass, "nchunks": nchunks,
                             "chunk_bytes": chunk_bytes, "xs": xs})
json.dump(doc, open(os.path.join(ROOT, "tests", "golden", "datapath_vectors.json"), "w"), indent=1)
print("wrote tests/golden/datapath_vectors.json")
