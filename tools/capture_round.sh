#!/bin/bash
# One GPU-box call that regenerates everything under profiles/ for a round tag:  bash tools/capture_round.sh r1j
# (each ncu pass runs only after the same command has exited 0 without ncu; numbers printed under ncu are never quoted)
TAG=${1:-rX}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -q 2>&1 | tail -4 > $O/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" >> $O/${TAG}_smoke.log
python bench.py --kernel-table $O/${TAG}_kernel_table.json > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err || exit 1
python bench.py --steps 1 --no-train --no-cpu-baseline > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file $O/${TAG}_launches_ncu.csv \
    python bench.py --steps 1 --no-train --no-cpu-baseline > $O/ncu_fwd.log 2>&1
python tools/bench_train.py --steps 1 > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $O/${TAG}_train_launches_ncu.csv \
    python tools/bench_train.py --steps 1 > $O/ncu_train.log 2>&1
python tools/fused_once.py > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:fused_ln_proj_dw_kernel -c 1 -f -o $O/${TAG}_fused \
    python tools/fused_once.py > $O/ncu_fused.log 2>&1
python tools/bench_ops.py gate k4 --once && {
ncu --set full --clock-control none --import-source on -k regex:dwconv_tma_kernel -c 1 -f -o $O/${TAG}_gate \
    python tools/bench_ops.py gate --once > $O/ncu_gate.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -c 1 -f -o $O/${TAG}_k4 \
    python tools/bench_ops.py k4 --once > $O/ncu_k4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_stream_tc_kernel -c 1 -f -o $O/${TAG}_k4_l4 \
    python tools/bench_ops.py k4 --levels l4 --once > $O/ncu_k4l4.log 2>&1
}
cat $O/${TAG}_pytest.log; tail -c 600 $O/${TAG}_bench.json
