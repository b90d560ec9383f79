#!/bin/bash
# round-2 GPU call M: GPU suite + bench on the final build (vector-load width chosen per tiling)
cd "$(dirname "$0")/.." || exit 1
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:$LD_LIBRARY_PATH
O=gpurun_out; T=${1:-r2m}
mkdir -p $O
rm -f $O/parity_tests.jsonl $O/closed_loop_tests.jsonl
timeout 2700 python -m pytest tests -m gpu -q > $O/${T}_pytest.log 2>&1
echo "pytest rc $?" >> $O/${T}_pytest.log
timeout 900 python bench.py > $O/${T}_bench_n1.json 2> $O/${T}_bench.err
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/${T}_smoke.log 2>&1
echo done
