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


def build(force=False, verbose=False):
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
    else:
        missing = [p for p in (f32, f64, base) if not os.path.exists(p)]
        if missing:
            raise RuntimeError(
                "reference checkout %s absent and prebuilt oracle/_ref artefacts missing: %s"
                % (REF, missing))
    return {"oracle": out, "qdldl_f32": f32, "qdldl_f64": f64, "cpu_baseline": base}


if __name__ == "__main__":
    print(build(force=True, verbose=True))
