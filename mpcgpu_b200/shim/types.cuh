// types.cuh -- drop-in for GBD-PCG's types.cuh: pcg_config<T> as used at
// include/mpcsim.cuh:250-253,261-262,269-270 and include/pcg/sqp.cuh:163-164 (the ADDRESSES of
// pcg_max_iter / pcg_exit_tol are passed as kernel arguments, so their types are the kernel's
// parameter types).
#pragma once
#include <cstdint>

template <typename T>
struct pcg_config {
    T pcg_exit_tol = static_cast<T>(1e-5);
    uint32_t pcg_max_iter = 200;
    unsigned pcg_grid = 0;      // unused: the caller launches grid = knot_points
    unsigned pcg_block = 128;
    bool pcg_org_trans = false; // unused (kept so aggregate users of the upstream struct compile)
    bool pcg_poly_order = false;
};
