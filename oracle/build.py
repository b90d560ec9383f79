"""TEST INFRASTRUCTURE ONLY -- build recipe for the CPU oracle.

Outputs
-------
oracle/_build/liboracle.so        our C restatement (mpc_oracle.c), fp32 + fp64
oracle/_ref/libqdldl_f32.so       the REFERENCE's qdldl/src/qdldl.c, float/int
                                  (configuration of the reference Makefile:16)
oracle/_ref/libqdldl_f64.so       same source, double/int (high-precision lambda*)
oracle/_ref/libcpu_baseline.so    cpu_baseline.c + mpc_oracle.c + reference qdldl.c (f32)

The `_ref` artefacts are compiled from the sources WHERE THEY LIE under
/root/reference (never copied into this repo).  /root/reference does not exist on the
GPU box: there the prebuilt `.so` files that travelled with the snapshot are used.
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("MPCGPU_REFERENCE", "/root/reference")
BUILD = os.path.join(HERE, "_build")
REFOUT = os.path.join(HERE, "_ref")
CFLAGS = ["-O2", "-ffp-contract=off", "-fPIC", "-shared", "-Wall"]


def _run(cmd):
    subprocess.run(cmd, check=True, cwd=HERE)


def _stale(out, srcs):
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.exists(s) and os.path.getmtime(s) > t for s in srcs)


NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
RUN = os.path.join(REFOUT, "run")            # working directory of the reference executables
HARNESS = os.path.join(RUN, "ref_harness")
TRACK = os.path.join(RUN, "track_iiwa_pcg_n%d")
TRACK_KNOTS = (32, 128)
# the same unchanged source with the SQP wall-clock budget lifted (-DSQP_MAX_TIME_US, an #ifndef knob of the reference's
# settings.cuh:173): under ncu every launch costs milliseconds and the stock 2 ms budget (even a 20 ms one) ends each SQP
# solve before its PCG launch, so the profiler's launch list is taken from this build, on a shortened copy of the
# trajectory files (tools/gpu_r2i.sh) so that the unprofiled run it must follow stays within minutes
TRACK_NOLIMIT = os.path.join(RUN, "track_iiwa_pcg_n128_nolimit")
SQPB = os.path.join(RUN, "sqp_batched_check")   # tools/sqp_batched_check.cu
KKTB = os.path.join(RUN, "kkt_batched_check")   # tools/kkt_batched_check.cu
SLOT = os.path.join(RUN, "track_slot_%s_n%d")    # tools/track_slot.cu: simulateMPC over the pcg / b200 slot
SLOTAB = os.path.join(RUN, "slot_ab_n%d")        # tools/slot_ab.cu: one SQP iteration of either slot on identical state
# product-side executable that needs the reference's dynamics headers to build: tools/e2e_producer.cu (batched KKT
# producer -> fused solve from host buffers).  NOT test infrastructure, so it is kept outside oracle/.
PLUGINS = os.path.join(os.path.dirname(HERE), "mpcgpu_b200", "_plugins")
E2EP = os.path.join(PLUGINS, "e2e_producer")


def _ref_nvcc(src, out, defs=(), link_product=False):
    root = os.path.dirname(HERE)
    inc = [os.path.join(root, "mpcgpu_b200", "shim"), os.path.join(root, "include"),
           os.path.join(REF, "include"), os.path.join(REF, "include", "common"),
           os.path.join(REF, "include", "pcg"), os.path.join(REF, "include", "dynamics"),
           os.path.join(REF, "include", "utils"), os.path.join(REF, "qdldl", "include"),
           os.path.join(HERE, "qdldl_cfg", "f32"), os.path.join(HERE, "qdldl_cfg")]
    cmd = [NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-w", "-lineinfo"]
    cmd += list(defs) + [a for i in inc for a in ("-I", i)] + [src, "-o", out, "-lcublas"]
    if link_product:   # the L2 slot calls the C ABI of mpcgpu_b200/libmpcgpu_b200.so
        rpath = "$ORIGIN/.." if link_product == "plugin" else "$ORIGIN/../../../mpcgpu_b200"
        cmd += ["-L", os.path.join(root, "mpcgpu_b200"), "-lmpcgpu_b200", "-Xlinker", "-rpath=" + rpath]
    _run(cmd)


def _build_reference_cuda(force, verbose):
    """The reference's OWN CUDA code (its form_schur_system kernel, KKT generation with the
    GRiD dynamics, and the whole examples/track_iiwa_pcg.cu tracking loop), compiled UNCHANGED
    from /root/reference against the shim headers (mpcgpu_b200/shim stands in for the empty
    GBD-PCG / GLASS submodules).  Outputs only under oracle/_ref/run/, which also receives a
    copy of the two trajectory fixtures the executables read by relative path."""
    if not os.path.exists(NVCC):
        return
    import glob
    import shutil
    root = os.path.dirname(HERE)
    deps = glob.glob(os.path.join(root, "mpcgpu_b200", "shim", "*.cuh")) + \
        glob.glob(os.path.join(root, "mpcgpu_b200", "csrc", "*.cuh")) + [os.path.join(HERE, "ref_harness.cu")]
    os.makedirs(os.path.join(RUN, "examples", "trajfiles"), exist_ok=True)
    os.makedirs(os.path.join(RUN, "tmp", "results"), exist_ok=True)
    for f in ("0_0_eepos.traj", "0_0_traj.csv"):
        dst = os.path.join(RUN, "examples", "trajfiles", f)
        if not os.path.exists(dst):
            shutil.copyfile(os.path.join(REF, "examples", "trajfiles", f), dst)
    jobs = []   # (src, out, defs, link_product)
    if force or _stale(HARNESS, deps):
        jobs.append((os.path.join(HERE, "ref_harness.cu"), HARNESS, [], False))
    for n in TRACK_KNOTS:
        if force or _stale(TRACK % n, deps):
            jobs.append((os.path.join(REF, "examples", "track_iiwa_pcg.cu"), TRACK % n, ["-DKNOT_POINTS=%d" % n], False))
    if force or _stale(TRACK_NOLIMIT, deps):
        jobs.append((os.path.join(REF, "examples", "track_iiwa_pcg.cu"), TRACK_NOLIMIT,
                     ["-DKNOT_POINTS=128", "-DSQP_MAX_TIME_US=100000000"], False))
    slot_src = os.path.join(root, "tools", "track_slot.cu")
    slot_deps = deps + [slot_src, os.path.join(root, "mpcgpu_b200", "shim", "b200", "sqp.cuh"),
                        os.path.join(root, "include", "mpcgpu_b200.h")]
    lib = os.path.join(root, "mpcgpu_b200", "libmpcgpu_b200.so")
    for n in TRACK_KNOTS:
        # pcg: the reference's slot on the whole L1 shim (our Schur + our PCG kernels); pcgrefschur: the same with the
        # reference's own form_schur_system kernel kept (the round-1 baseline); b200: the L2 slot
        for name, defs in (("pcg", ["-DSLOT_B200=0"]), ("pcgrefschur", ["-DSLOT_B200=0", "-DMPCGPU_B200_SHIM_KEEP_REFERENCE_SCHUR=1"]),
                           ("b200", ["-DSLOT_B200=1"])):
            flag = name == "b200"
            if flag and not os.path.exists(lib):
                continue   # product library not built yet (mpcgpu_b200.build runs first in __graft_entry__.build)
            if force or _stale(SLOT % (name, n), slot_deps):
                jobs.append((slot_src, SLOT % (name, n), ["-DKNOT_POINTS=%d" % n] + defs, flag))
    ab_src = os.path.join(root, "tools", "slot_ab.cu")
    for n in TRACK_KNOTS:
        if os.path.exists(lib) and (force or _stale(SLOTAB % n, slot_deps + [ab_src])):
            jobs.append((ab_src, SLOTAB % n, ["-DKNOT_POINTS=%d" % n, "-DSQP_MAX_TIME_US=100000000"], True))
    ep_src = os.path.join(root, "tools", "e2e_producer.cu")
    ep_deps = deps + [ep_src, os.path.join(root, "mpcgpu_b200", "shim", "b200", "kkt_batched.cuh"), os.path.join(root, "include", "mpcgpu_b200.h")]
    if os.path.exists(lib) and (force or _stale(E2EP, ep_deps)):
        os.makedirs(os.path.join(PLUGINS, "examples", "trajfiles"), exist_ok=True)
        for f in ("0_0_eepos.traj", "0_0_traj.csv"):
            dst = os.path.join(PLUGINS, "examples", "trajfiles", f)
            if not os.path.exists(dst):
                shutil.copyfile(os.path.join(REF, "examples", "trajfiles", f), dst)
        jobs.append((ep_src, E2EP, [], "plugin"))
    kb_src = os.path.join(root, "tools", "kkt_batched_check.cu")     # batched KKT producer vs the reference kernel
    kb_deps = deps + [kb_src, os.path.join(root, "mpcgpu_b200", "shim", "b200", "kkt_batched.cuh"),
                      os.path.join(root, "include", "mpcgpu_b200.h")]
    if os.path.exists(lib) and (force or _stale(KKTB, kb_deps)):
        jobs.append((kb_src, KKTB, [], True))
    sb_src = os.path.join(root, "tools", "sqp_batched_check.cu")     # batched SQP iterations vs the reference's sequence
    sb_deps = kb_deps + [sb_src, os.path.join(root, "mpcgpu_b200", "shim", "b200", "sqp_batched.cuh")]
    if os.path.exists(lib) and (force or _stale(SQPB, sb_deps)):
        jobs.append((sb_src, SQPB, [], True))
    if jobs:   # each is a minute of nvcc on the GRiD dynamics headers: run them side by side
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=min(len(jobs), max(1, (os.cpu_count() or 2) // 2))) as ex:
            for _ in ex.map(lambda j: _ref_nvcc(j[0], j[1], j[2], link_product=j[3]), jobs):
                pass
        if verbose:
            print("built", *[j[1] for j in jobs])


def build(force=False, verbose=False, reference_cuda=True):
    """reference_cuda=False (what `import oracle` uses) builds only the C libraries -- seconds; the
    reference's CUDA executables (minutes of nvcc) are built by __graft_entry__.build() / `python -m oracle.build`."""
    os.makedirs(BUILD, exist_ok=True)
    os.makedirs(REFOUT, exist_ok=True)
    own = [os.path.join(HERE, f) for f in ("mpc_oracle.c", "oracle_impl.h")]
    out = os.path.join(BUILD, "liboracle.so")
    if force or _stale(out, own):
        _run(["gcc", *CFLAGS, "-o", out, "mpc_oracle.c", "-lm"])
        if verbose:
            print("built", out)

    qsrc = os.path.join(REF, "qdldl", "src", "qdldl.c")
    qinc = os.path.join(REF, "qdldl", "include")
    have_ref = os.path.exists(qsrc)
    f32 = os.path.join(REFOUT, "libqdldl_f32.so")
    f64 = os.path.join(REFOUT, "libqdldl_f64.so")
    base = os.path.join(REFOUT, "libcpu_baseline.so")
    if have_ref:
        # qdldl_types.h / qdldl_version.h are cmake-generated in the reference
        # (qdldl/configure/*.in); oracle/qdldl_cfg holds the two configurations we build.
        cfg = os.path.join(HERE, "qdldl_cfg")
        i32 = ["-I", os.path.join(cfg, "f32"), "-I", cfg, "-I", qinc]
        i64 = ["-I", os.path.join(cfg, "f64"), "-I", cfg, "-I", qinc]
        if force or _stale(f32, [qsrc]):
            _run(["gcc", "-O3", "-fPIC", "-shared", *i32, "-o", f32, qsrc])
        if force or _stale(f64, [qsrc]):
            _run(["gcc", "-O3", "-fPIC", "-shared", *i64, "-o", f64, qsrc])
        if force or _stale(base, own + [qsrc, os.path.join(HERE, "cpu_baseline.c")]):
            _run(["gcc", "-O3", "-ffp-contract=off", "-fPIC", "-shared", "-pthread",
                  *i32, "-o", base, "cpu_baseline.c", "mpc_oracle.c", qsrc, "-lm"])
        if verbose:
            print("built", f32, f64, base)
        if reference_cuda:
            _build_reference_cuda(force, verbose)
    else:
        missing = [p for p in (f32, f64, base) if not os.path.exists(p)]
        if missing:
            raise RuntimeError(
                "reference checkout %s absent and prebuilt oracle/_ref artefacts missing: %s"
                % (REF, missing))
    return {"oracle": out, "qdldl_f32": f32, "qdldl_f64": f64, "cpu_baseline": base}


if __name__ == "__main__":
    print(build(force=True, verbose=True))
