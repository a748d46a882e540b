"""Writes tests/golden/fold_ref_tables.json: the miss-rate tables the REFERENCE binaries print
(oracle/_ref/barrett_final = rfhe_framewk/src/barrett_final.cpp, oracle/_ref/outer_check =
rfhe_framewk/src/barrett_test/outer_check.cpp, both built by oracle/Makefile from the sources where they lie).
Run in the authoring container (the binaries do not travel with the history): python tests/golden/make_fold_golden.py"""
import json
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = os.path.join(ROOT, "oracle", "_ref")
out = {"barrett_final": {}, "outer_check": {}}
VEC, TRIALS, FOLDS = 64, 200000, "1,2,8,64"
for mode in ["SBF", "DBF", "SBF+SBF", "SBF+DBF", "MOF1", "MOF2"]:
    txt = subprocess.run([os.path.join(REF, "barrett_final"), "--mode", mode, "--trials", str(TRIALS), "--vec-len", str(VEC),
                          "--folds", FOLDS, "--seed", "1"], capture_output=True, text=True, check=True).stdout
    hdr = txt.splitlines()[0]
    for line in txt.splitlines():
        p = line.split(",")
        if len(p) == 6 and p[0] in ("instra", "sum"):
            out["barrett_final"]["%s/%s/%s" % (mode, p[0], p[1])] = {"trials": int(p[2]), "missed": int(p[4])}
    out["barrett_final"]["header"] = hdr
out["barrett_final"]["args"] = {"vec_len": VEC, "trials": TRIALS, "folds": FOLDS, "seed": 1}
N, QB, TR = 64, 37, 400000
txt = subprocess.run([os.path.join(REF, "outer_check"), "--n", str(N), "--qbits", str(QB), "--trials", str(TR), "--fold_min", "4",
                      "--fold_max", "16", "--fold_step", "4", "--modes", "SBF,DBF,SBF+SBF,SBF+DBF,MOF1,MOF2"],
                     capture_output=True, text=True, check=True).stdout
for line in txt.splitlines()[1:]:
    strat, mode, k, rate = line.split(",")
    out["outer_check"]["%s/%s/%s" % (mode, strat, k)] = {"trials": TR, "missed": round(float(rate) * TR)}
out["outer_check"]["args"] = {"n": N, "qbits": QB, "trials": TR}
with open(os.path.join(ROOT, "tests", "golden", "fold_ref_tables.json"), "w") as f:
    json.dump(out, f, indent=1, sort_keys=True)
print("wrote fold_ref_tables.json:", len(out["barrett_final"]), "+", len(out["outer_check"]), "entries")
