"""Writes profiles/sass/<kernel>.sass for the hot kernels of tiny-cuda-nn_b200/lib/libtcnn_b200.so and INDEX.md with the counts of the
mnemonics that prove the tensor-core / TMA / reduction paths (B200_PROFILING.md: UTCHMMA = tcgen05.mma kind::f16, LDTM = tcgen05.ld,
UTCBAR = tcgen05.commit, UBLKCP = cp.async.bulk, REDG = red.global, SYNCS = mbarrier, LDGMC = multimem.ld_reduce).

    python profiles/sass/extract_sass.py            # needs cuobjdump + c++filt; no GPU
"""
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
LIB = os.path.join(ROOT, "tiny-cuda-nn_b200", "lib", "libtcnn_b200.so")

# file name -> demangled prefix of the kernel (first match)
KERNELS = {
    "mlp_fused_train_64x2_grid2d_f2": "void b200::mlp_kernel<1, 2, 2, 2, 0, false, false>(",       # metric config: fwd + loss + bwd + dW + grid scatter
    "mlp_fused_train_64x2_grid3d_f2": "void b200::mlp_kernel<1, 2, 3, 2, 0, false, false>(",
    "mlp_inference_64x2": "void b200::mlp_kernel<0, 2, 0, 2, 0, false, false>(",
    "grid_forward_2d_f2": "void b200::grid_forward_kernel<2, 2>(",
    "grid_forward_3d_f2": "void b200::grid_forward_kernel<3, 2>(",
    "grid_backward_2d_f2": "void b200::grid_backward_kernel<2, 2>(",
    "adam": "b200::adam_kernel(",
    "dp_adam_multicast": "void b200::dp_adam_kernel<true, 8>(",
    "dp_adam_peer": "void b200::dp_adam_kernel<false, 8>(",
    "mlp_finalize": "b200::mlp_finalize_kernel(",
    "mlp_finalize_adam": "b200::mlp_finalize_adam_kernel(",
    "mlp_wide_forward_train": "void b200::mlp_wide_forward_kernel<1, 1>(",
    "mlp_wide_backward": "void b200::mlp_wide_backward_kernel<1>(",
    "mlp_wide_dw": "b200::mlp_wide_dw_kernel(",
    "oneblob_encode": "b200::oneblob_encode_kernel(",
}
MNEMONICS = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UBLKCP", "UTMALDG", "SYNCS", "REDG", "ATOMG", "LDGMC", "LDG", "STG", "LDS", "STS", "STL", "LDL", "BAR.SYNC", "HFMA2", "FFMA", "IMAD"]


def main():
    text = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    parts = re.split(r"^\s*Function : ", text, flags=re.M)[1:]
    funcs = {}
    for p in parts:
        mangled, body = p.split("\n", 1)
        name = subprocess.run(["c++filt", mangled.strip()], capture_output=True, text=True).stdout.strip()
        funcs[name] = (mangled.strip(), body)
    rows = []
    for fname, prefix in KERNELS.items():
        hit = [n for n in funcs if n.startswith(prefix)]
        if not hit:
            print("not found:", prefix, file=sys.stderr)
            continue
        name = hit[0]
        mangled, body = funcs[name]
        lines = [l for l in body.splitlines() if re.search(r"/\*[0-9a-f]{4}\*/", l)]
        instr = [re.sub(r"^\s*/\*[0-9a-f]+\*/\s*", "", l).split("/*")[0].strip() for l in lines]
        # keep mnemonic text only (drop the encoding columns): the listing stays small enough to read and to diff
        with open(os.path.join(HERE, fname + ".sass"), "w") as f:
            f.write(f"// {name}\n// {mangled}\n// cuobjdump -sass of tiny-cuda-nn_b200/lib/libtcnn_b200.so (sm_100a), encodings stripped\n")
            f.write("\n".join(instr) + "\n")
        counts = {m: sum(1 for i in instr if re.search(r"(^|\s)" + re.escape(m), re.sub(r"^@!?U?P\d+\s+", "", i))) for m in MNEMONICS}
        red_kinds = sorted({re.sub(r"^@!?U?P\d+\s+", "", i).split()[0] for i in instr if "REDG" in i or "LDGMC" in i or "MMEM" in i})
        rows.append((fname, name, len(instr), counts, red_kinds))
    with open(os.path.join(HERE, "INDEX.md"), "w") as f:
        f.write("# SASS of the hot kernels (sm_100a)\n\nWritten by `python profiles/sass/extract_sass.py` from the library `__graft_entry__.build()` produces "
                "(`nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo`).\nOne `.sass` file per kernel (instruction text, encodings stripped).  "
                "Counts are static instruction counts in the listing, not executed counts.\n\n")
        f.write("| file | kernel | instr | " + " | ".join(MNEMONICS) + " | reduction forms |\n|---|---|---|" + "---|" * (len(MNEMONICS) + 1) + "\n")
        for fname, name, n, c, rk in rows:
            f.write(f"| `{fname}.sass` | `{name.split('(')[0]}` | {n} | " + " | ".join(str(c[m]) for m in MNEMONICS) + " | " + ", ".join(rk) + " |\n")
        f.write("\nReading the table: `UTCHMMA` = `tcgen05.mma.cta_group::1.kind::f16` (the MMAs of the layer chain and of the weight gradients), "
                "`LDTM` = `tcgen05.ld` (…`.PACK16BIT` = the packed fp16-accumulator read), `UTCBAR` = `tcgen05.commit` onto an mbarrier, "
                "`UBLKCP` = `cp.async.bulk` (weights shared->global / global->shared), `SYNCS` = mbarrier arrive / try_wait, "
                "`REDG.E.ADD.F16x2...` = the grid-gradient reductions (`red.global.add.noftz.f16x2` and the `.v2/.v4` forms), `LDGMC...HPADD.F16x4` = `multimem.ld_reduce` (NVSwitch in-fabric sum of the ranks' gradients) of the "
                "data-parallel optimizer.  `STL`/`LDL` = 0 means no register spills.\n")
    for r in rows:
        print(r[0], r[2], {k: v for k, v in r[3].items() if v}, r[4])


if __name__ == "__main__":
    main()
