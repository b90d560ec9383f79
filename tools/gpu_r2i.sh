#!/bin/bash
# round-2 GPU call I: drift of the PCG variants against the exact-arithmetic iterate (default library = variant 0 vs the variant-2 experiment build libexp_m2.so);
# ncu launch list of the reference's unchanged closed loop (N = 128) on the L1 shim, on a shortened trajectory copy
cd "$(dirname "$0")/.." || exit 1
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:$LD_LIBRARY_PATH
O=$PWD/gpurun_out; T=${1:-r2i}
mkdir -p $O
# libexp_m2.so: tools/ab_build.sh m2 -DMPCB200_PCG_VARIANT=2 (the run-time opt-in covers the fused kernel only; this tool drives the PCG kernel)
timeout 900 python tools/variant_drift.py default m2 > $O/${T}_variant_drift.jsonl 2> $O/${T}_variant_drift.err
R=oracle/_ref/run
mkdir -p $R/short/examples/trajfiles $R/short/tmp/results
head -n 136 $R/examples/trajfiles/0_0_eepos.traj > $R/short/examples/trajfiles/0_0_eepos.traj
head -n 136 $R/examples/trajfiles/0_0_traj.csv > $R/short/examples/trajfiles/0_0_traj.csv
( cd $R/short
  timeout 280 ../track_iiwa_pcg_n128_nolimit > $O/${T}_l1_plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 150 --csv --log-file $O/${T}_l1_track_n128_launches.csv ../track_iiwa_pcg_n128_nolimit > $O/${T}_l1_ncu.log 2>&1 )
echo done
