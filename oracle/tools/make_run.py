#!/usr/bin/env python
"""Set up a run directory for the shim-built reference (TEST INFRASTRUCTURE; needs /root/reference).

  python oracle/tools/make_run.py classic1d oracle/_ref/runs/s1 [--ftstep 20] [--nlevels 100] ...

Mirrors what the reference's tests/setup_classicmode.sh does (rsync the input files, cp ../data/*), with the
network-hosted atomic data replaced by oracle/tools/gen_atomic_data.py. Everything is written under oracle/_ref/
(git-ignored); nothing from the reference is committed to this repository.
"""
import argparse
import os
import shutil
import sys

sys.path.insert(0, os.path.dirname(__file__))
import gen_atomic_data  # noqa: E402

REF = os.environ.get("ARTIS_REF", "/root/reference")


def write_input_txt(path, template, itstep, ftstep, ntstep=None, tmin=None, tmax=None, model_type=None, ngrey=None,
                    resume=False):
    lines = open(template).read().splitlines()
    out = []
    k = 0
    for line in lines:
        if line.strip().startswith("#") or not line.strip():
            out.append(line)
            continue
        body, _, comment = line.partition("#")
        if k == 1 and ntstep is not None:
            body = f"{ntstep}".ljust(25)
        if k == 2:
            body = f"{itstep:03d} {ftstep:03d}".ljust(25)
        if k == 3 and tmin is not None:
            body = f"{tmin} {tmax}".ljust(25)
        if k == 7 and model_type is not None:
            body = f"{model_type}".ljust(25)
        if k == 16 and resume:  # simulation_continued_from_saved
            body = "1".ljust(25)
        if k == 19 and ngrey is not None:  # cell_is_optically_thick num_grey_timesteps
            body = f"{body.split()[0]} {ngrey}".ljust(25)
        out.append(body + ("#" + comment if comment else ""))
        k += 1
    open(path, "w").write("\n".join(out) + "\n")


# SURVEY 8d S5 (BASELINE.json config 5): five observers at phi = 0, one spectrum each (all opacities), the compiled time and
# frequency windows, the input file's thick-cell limit, tau_max 10, no velocity maps
VPKT_TXT_S5 = """5
-0.8 -0.4 0 0.4 0.8
0 0 0 0 0
0
0 0 0
0
0 0
10
0
"""

VPKT_TXT = """2
0.5 -0.3
0
90
1 6 0 26 -1 -2 -3 -4
0 0 0
0
0 0
10
1
10 30
2 3500 6000 6000 10000
"""


def setup_classic1d(rundir, args):
    src = os.path.join(REF, "tests", "classicmode_inputfiles")
    os.makedirs(rundir, exist_ok=True)
    for f in ("model.txt", "abundances.txt"):
        shutil.copy(os.path.join(src, f), rundir)
    for f in os.listdir(os.path.join(REF, "data")):
        shutil.copy(os.path.join(REF, "data", f), rundir)
    write_input_txt(os.path.join(rundir, "input.txt"), os.path.join(src, "input-newrun.txt"), args.itstep, args.ftstep,
                    resume=args.resume)
    if args.vpkt:
        # vpkt.txt (vpkt.cc:667-836): two observers; six spectra per observer (all opacities, iron lines removed, no
        # lines, no bf, no ff, no es); compiled time and frequency windows; tau_max 10; velocity maps in two bands
        open(os.path.join(rundir, "vpkt.txt"), "w").write(VPKT_TXT_S5 if args.vpkt == "s5" else VPKT_TXT)
    n = gen_atomic_data.generate(rundir, tuple(int(z) for z in args.elements.split(",")), args.nions, args.nlevels,
                                 args.dense_levels, args.sparsity, args.seed)
    print(f"run dir {rundir}: {n} lines")


def setup_kilonova1d(rundir, args):
    """tests/setup_kilonova.sh: the 1D kilonova model with the composition of tests/kilonova_inputfiles (Fe I-V,
    Co II-IV, Ni II-V with levels and lines, seven abundance-only elements), synthetic atomic data"""
    import lzma
    src = os.path.join(REF, "tests", "kilonova_inputfiles")
    os.makedirs(rundir, exist_ok=True)
    shutil.copy(os.path.join(src, "abundances.txt"), rundir)
    for f in ("model.txt", "gridcontributions.txt"):
        open(os.path.join(rundir, f), "wb").write(lzma.open(os.path.join(src, f + ".xz")).read())
    for f in os.listdir(os.path.join(REF, "data")):
        shutil.copy(os.path.join(REF, "data", f), rundir)
    write_input_txt(os.path.join(rundir, "input.txt"), os.path.join(src, "input-newrun.txt"), args.itstep, args.ftstep,
                    ngrey=args.ngrey, resume=args.resume)
    n = gen_atomic_data.generate(rundir, (26, 27, 28), 5, args.nlevels, args.dense_levels, args.sparsity, args.seed,
                                 ion_ranges={26: (1, 5), 27: (2, 4), 28: (2, 5)},
                                 abundance_only=[(2, 4), (24, 60), (38, 95), (58, 145), (60, 150), (88, 220), (92, 230)],
                                 single_level_top_ion=False)
    print(f"run dir {rundir}: {n} lines")


def setup_nebular1zone(rundir, args):
    """tests/setup_nebularonezone.sh: the one-zone nebular model (model.txt, abundances.txt, recombrates.txt,
    input-newrun.txt of tests/nebularonezone_inputfiles verbatim) with synthetic Fe/Co/Ni atomic data"""
    src = os.path.join(REF, "tests", "nebularonezone_inputfiles")
    os.makedirs(rundir, exist_ok=True)
    for f in ("model.txt", "abundances.txt", "recombrates.txt"):
        shutil.copy(os.path.join(src, f), rundir)
    for f in os.listdir(os.path.join(REF, "data")):
        shutil.copy(os.path.join(REF, "data", f), rundir)
    write_input_txt(os.path.join(rundir, "input.txt"), os.path.join(src, "input-newrun.txt"), args.itstep, args.ftstep,
                    resume=args.resume)
    n = gen_atomic_data.generate(rundir, (26, 27, 28), args.nions, args.nlevels, args.dense_levels, args.sparsity,
                                 args.seed, single_level_top_ion=False)
    print(f"run dir {rundir}: {n} lines")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("kind", choices=["classic1d", "kilonova1d", "nebular1zone"])
    ap.add_argument("rundir")
    ap.add_argument("--itstep", type=int, default=0)
    ap.add_argument("--ftstep", type=int, default=36)
    ap.add_argument("--elements", default="14,16,20,26,27,28")
    ap.add_argument("--nions", type=int, default=5)
    ap.add_argument("--nlevels", type=int, default=60)
    ap.add_argument("--dense-levels", type=int, default=15)
    ap.add_argument("--sparsity", type=float, default=0.15)
    ap.add_argument("--seed", type=int, default=20240607)
    ap.add_argument("--ngrey", type=int, default=None, help="kilonova1d: override num_grey_timesteps of input.txt")
    ap.add_argument("--vpkt", nargs="?", const="fixture", default=None,
                    help="classic1d: also write a vpkt.txt for VPKT_ON builds (`--vpkt s5`: the five observers of config 5)")
    ap.add_argument("--resume", action="store_true",
                    help="write input.txt to continue from the gridsave/packets files of timestep itstep-1")
    args = ap.parse_args()
    if args.kind == "classic1d":
        setup_classic1d(args.rundir, args)
    elif args.kind == "kilonova1d":
        setup_kilonova1d(args.rundir, args)
    else:
        setup_nebular1zone(args.rundir, args)


if __name__ == "__main__":
    main()
