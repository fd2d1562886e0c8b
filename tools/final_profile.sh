#!/bin/bash
# Round-end measurement on ONE GPU (run under gpurun): the bench line, the ncu launch list of the same short bench command,
# one full capture of the prefill kernel (C2) and one of the decode kernel (C3), the shape sweeps.  Outputs in gpurun_out/.
# Every ncu command is preceded by the same command without ncu (the recipe in B200_PROFILING.md).
R=${1:-r02}
O=gpurun_out
SHORT="python bench.py --steps 2 --warmup 1 --no-sharded --no-ref-kernel --no-cpu --no-sustained"
python bench.py > $O/${R}_bench_line.json 2> $O/${R}_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > $O/${R}_bench_reference_arm.json 2>> $O/${R}_bench.err
$SHORT > $O/plain1.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${R}_launches_bench.csv $SHORT > $O/ncu_l.log 2>&1
$SHORT --no-decode > $O/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fa_prefill_tc5 -s 1 -c 1 -o $O/${R}_prof_prefill -f $SHORT --no-decode > $O/ncu_p.log 2>&1
python tools/c3_once.py > $O/plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fa_decode -s 2 -c 1 -o $O/${R}_prof_decode -f python tools/c3_once.py > $O/ncu_d.log 2>&1
python tools/shape_bench.py > $O/${R}_shapes.jsonl 2> $O/shapes.err
python tools/prefill_bench.py --iters 30 2>/dev/null | grep config > $O/${R}_prefill_shapes.jsonl
python tools/c4_shard_probe.py > $O/${R}_c4_shard_probe.jsonl 2>> $O/shapes.err
python tools/c1_latency.py > $O/${R}_c1_latency.jsonl 2>> $O/shapes.err
tail -c 600 $O/${R}_bench_line.json; echo; tail -3 $O/ncu_p.log; tail -3 $O/ncu_d.log
