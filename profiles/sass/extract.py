"""Per-kernel SASS instruction histograms of lib/libb200fhe.so (cuobjdump -sass), written to profiles/sass/<name>.txt:
mnemonic counts (proof of which units the hot kernels use: IMAD.WIDE chains, IMMA.16832.U8.U8, UBLKCP = TMA bulk copy,
SYNCS = mbarrier, REDUX, ...) plus the first 60 instructions.  python profiles/sass/extract.py"""
import collections, os, re, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
LIB = os.path.join(ROOT, "fhe_reliability_gpu_b200", "lib", "libb200fhe.so")
WANT = {"k_fwd_col": "k_fwd_colILi8ELi4ELi2ELb0ELb0", "k_fwd_col_prot": "k_fwd_colILi8ELi4ELi2ELb1ELb0", "k_fwd_row": "k_fwd_rowILi2ELb0ELb0",
        "k_fwd_row_prot": "k_fwd_rowILi2ELb1ELb0", "k_inv_row": "k_inv_rowILi2E", "k_inv_col": "k_inv_colILi8ELi4ELi2E",
        "k_bconv_contract_tc": "k_bconv_contract_tcILi4ELb0", "k_bconv_contract_tc5": "k_bconv_contract_tc5", "k_pmult": "k_pmult", "k_barrett_fold_checks": "k_barrett_foldILb1ELb1"}
txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)
for name, key in WANT.items():
    body = next((f for f in funcs if f.startswith("_ZN") and key in f.split("\n", 1)[0]), None)
    if body is None:
        continue
    ins = re.findall(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", body)
    cnt = collections.Counter(ins)
    with open(os.path.join(ROOT, "profiles", "sass", name + ".txt"), "w") as f:
        f.write("# %s\n# %d instructions; mnemonic histogram (cuobjdump -sass, sm_100a)\n" % (body.split("\n", 1)[0], len(ins)))
        for k, v in cnt.most_common():
            f.write("%6d  %s\n" % (v, k))
        f.write("\n# first 60 instructions\n")
        lines = [l for l in body.split("\n") if re.search(r"/\*[0-9a-f]{4}\*/", l)]
        f.write("\n".join(l.rstrip()[:120] for l in lines[:60]) + "\n")
    print(name, len(ins), {k: cnt[k] for k in cnt if k.startswith(("IMAD.WIDE", "IMAD.HI", "IMMA", "UBLKCP", "SYNCS", "REDUX", "UTMA", "LDGSTS", "UTC", "LDTM", "DFMA", "DADD"))})
