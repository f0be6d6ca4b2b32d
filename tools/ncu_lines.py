#!/usr/bin/env python
"""Attribute an `ncu --page source --print-source sass --csv` export to source lines through the inline chains that
`nvdisasm -gi` prints for the same cubin.

usage: ncu_lines.py SASS_CSV NVDISASM_TXT KERNEL_SUBSTRING [topN]
Prints, for the kernel, the top source lines by warp instructions executed (exclusive = innermost frame, inclusive = any
frame of the inline chain) with their thread instructions, average active lanes and stall samples.
"""
import csv, re, sys, collections

def parse_disasm(path, kernel):
    chains, cur, pending, inside = {}, [], [], False
    for ln in open(path, errors="replace"):
        if ln.startswith("//---") and ".text." in ln:
            inside = kernel in ln
            continue
        if not inside:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
        if m:
            pending.append((m.group(1).split("/")[-1], int(m.group(2))))
            continue
        m = re.match(r"\s*/\*([0-9a-f]+)\*/", ln)
        if m:
            if pending:
                cur, pending = pending, []
            chains[int(m.group(1), 16)] = cur
    return chains

def main():
    sass_csv, dis, kernel = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    chains = parse_disasm(dis, kernel)
    rows = list(csv.reader(open(sass_csv)))
    hdr = rows[1]
    ia, ii, it, ismp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
    # stall reasons of the sampled warps (columns of the source page); the first is the sort key of the long_sb% column
    stall_names = [n for n in ("stall_long_sb", "stall_no_inst", "stall_wait", "stall_short_sb", "stall_imc_miss", "stall_lg_throttle",
                               "stall_mio_throttle", "stall_math_throttle", "stall_branch_resolving", "stall_barrier", "stall_membar",
                               "stall_sleeping", "stall_dispatch") if n in hdr]
    cols = [hdr.index(n) for n in stall_names]
    nv = 3 + len(cols)
    base = None
    excl = collections.defaultdict(lambda: [0] * nv)
    incl = collections.defaultdict(lambda: [0] * nv)
    tot = [0] * nv
    for r in rows[2:]:
        if len(r) <= max(cols + [ismp]) or not r[ia].startswith("0x"):
            continue
        a = int(r[ia], 16)
        if base is None:
            base = a
        ch = chains.get(a - base, [("?", 0)])
        v = [int(r[ii] or 0), int(r[it] or 0), int(r[ismp] or 0)] + [int(r[c] or 0) for c in cols]
        for j in range(nv):
            tot[j] += v[j]
            excl[ch[0]][j] += v[j]
        for fr in set(ch):
            for j in range(nv):
                incl[fr][j] += v[j]
    print(f"total warp-inst {tot[0]:.3e} thread-inst {tot[1]:.3e} lanes/inst {tot[1]/max(tot[0],1):.2f} samples {tot[2]}")
    print("stall samples, % of all samples: " + "  ".join(f"{n[6:]} {100*tot[3+i]/max(tot[2],1):.1f}" for i, n in enumerate(stall_names)))
    short = [n[6:9] for n in stall_names]
    for name, d in (("EXCLUSIVE (innermost line)", excl), ("INCLUSIVE (any inline frame)", incl)):
        print("\n" + name + "   [stall columns: % of the LINE's samples]")
        print(f"{'file:line':34s} {'warp-inst%':>10s} {'lanes':>6s} {'samples%':>9s}  " + " ".join(f"{x:>4s}" for x in short))
        for k, v in sorted(d.items(), key=lambda kv: -kv[1][2])[:top]:
            print(f"{k[0]+':'+str(k[1]):34s} {100*v[0]/tot[0]:10.2f} {v[1]/max(v[0],1):6.2f} {100*v[2]/max(tot[2],1):9.2f}  " +
                  " ".join(f"{100*v[3+i]/max(v[2],1):4.0f}" for i in range(len(cols))))

if __name__ == "__main__":
    main()
