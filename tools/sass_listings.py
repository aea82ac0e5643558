"""Regenerate ``profiles/sass/``: one ``cuobjdump -sass`` listing per kernel of the built library (instruction
encodings stripped) plus a README with static opcode counts.

    python tools/sass_listings.py            # needs mpitree_b200/lib/libmpitree_b200.so (``__graft_entry__.build()``)

Template instantiations that only differ in the features-per-tile parameter are listed once (FT = 16).
"""
import collections
import pathlib
import re
import subprocess
import sys

ROOT = pathlib.Path(__file__).resolve().parents[1]
LIB = ROOT / "mpitree_b200" / "lib" / "libmpitree_b200.so"
OUT = ROOT / "profiles" / "sass"

OPS = ["UBLKCP", "UTMALDG", "SYNCS", "ATOMS", "IDP", "LDGMC", "REDG", "ATOMG", "RED", "LDG", "STG", "LDS", "STS", "LDL", "STL",
       "DADD", "DMUL", "DFMA", "MUFU", "PRMT", "BAR", "SHFL", "MEMBAR", "VOTE", "POPC", "UCGABAR", "CCTL"]

# mangled-name fragment -> file stem
WANT = [
    ("k_level_fused2ILi16ELb1ELb0E", "k_level_fused2_FT16_interleaved"),
    ("k_level_fused2ILi16ELb1ELb1E", "k_level_fused2_FT16_interleaved_nvls"),
    ("k_level_fused2ILi16ELb0ELb0E", "k_level_fused2_FT16_planar"),
    ("k_level_fusedILi16ELb1ELb0E", "k_level_fused_FT16_interleaved"),
    ("k_root_direct", "k_root_direct"),
    ("k_root_cluster", "k_root_cluster"),
    ("k_histILi16ELb1E", "k_hist_FT16_interleaved"),
    ("k_histILi16ELb0E", "k_hist_FT16_planar"),
    ("k_gainE", "k_gain"), ("k_finalize", "k_finalize"), ("k_plan", "k_plan"), ("k_partition", "k_partition"),
    ("k_commit", "k_commit"), ("k_push", "k_push"), ("k_klogk_eval", "k_klogk_eval"),
    ("k_encode_f32_slots", "k_encode_f32_slots"), ("k_encode_f32_quads", "k_encode_f32_quads"), ("k_encode_floatIfE", "k_encode_float_f32"),
    ("k_encode_u8", "k_encode_u8"), ("k_retile_u8", "k_retile_u8"),
    ("k_predict_u8_compact", "k_predict_u8_compact"), ("k_predict_stagedIffE", "k_predict_staged_f32"),
    ("k_subtreeILi8E", "k_subtree_FT8"), ("k_gather_rows", "k_gather_rows"),
    ("k_allreduce_oneshot", "k_allreduce_oneshot"), ("k_allreduce_twoshot", "k_allreduce_twoshot"),
    ("k_allreduce_nvls", "k_allreduce_nvls"),
]


def main():
    if not LIB.exists():
        sys.exit(f"{LIB} is missing: build first")
    text = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    funcs = collections.OrderedDict()
    name = None
    for line in text.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = m.group(1)
            funcs[name] = []
            continue
        if name is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(.*?)\s*;?\s*/\*", line)
        if m:
            funcs[name].append(m.group(1).rstrip(" ;"))
    OUT.mkdir(parents=True, exist_ok=True)
    for old in OUT.glob("*.sass"):
        old.unlink()
    rows = []
    for frag, stem in WANT:
        match = [n for n in funcs if frag in n]
        if not match:
            continue
        body = funcs[match[0]]
        (OUT / f"{stem}.sass").write_text(f"// {match[0]}\n" + "\n".join(body) + "\n")
        cnt = collections.Counter()
        for ins in body:
            op = ins.split()[1] if ins.startswith("@") and len(ins.split()) > 1 else ins.split()[0]
            base = op.split(".")[0]
            if base in OPS:
                cnt[base] += 1
        rows.append((stem, len(body), ", ".join(f"{k} x{cnt[k]}" for k in OPS if cnt[k])))
    with open(OUT / "README.md", "w") as fh:
        fh.write("# SASS listings (cuobjdump -sass of mpitree_b200/lib/libmpitree_b200.so, sm_100a; encodings stripped)\n\n")
        fh.write("Regenerate with `python tools/sass_listings.py`.  One file per kernel; static opcode counts of interest:\n\n")
        fh.write("| kernel | SASS instructions | notable opcodes |\n|---|---|---|\n")
        for stem, n, ops in rows:
            fh.write(f"| `{stem}` | {n} | {ops} |\n")
        fh.write("\n`UBLKCP` = `cp.async.bulk` (TMA bulk copy), `UTMALDG` = `cp.async.bulk.tensor` (tiled TMA load, the multicast copy of\n"
                 "`k_root_cluster`), `SYNCS` = mbarrier operations, `ATOMS` = shared-memory atomics (`red.shared.add` / `ATOMS.POPC.INC`),\n"
                 "`IDP` = `IDP.4A` (byte dot product: forms a histogram address - byte extract, scale, add - in one instruction),\n"
                 "`REDG` = `red.global.add`, `LDGMC` = `multimem.ld_reduce` (in-switch reduction), `UCGABAR` = cluster barrier;\n"
                 "`MEMBAR.SYS` + `ST.E.STRONG.SYS` / `LD.E.STRONG.SYS` form the peer barrier.  No tcgen05 / `UTC*MMA` by design: the\n"
                 "workload has no GEMM (DESIGN.md §2).\n")
    print(f"wrote {len(rows)} listings to {OUT}")


if __name__ == "__main__":
    main()
