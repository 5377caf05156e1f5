#!/bin/bash
# Round-end evidence run on the GPU box: parity tests, the full bench (both arms), the ncu launch list of the bench
# command and one `ncu --set full` capture of every kernel (second step of a 128-pair bench).  Outputs -> gpurun_out/.
TAG=${1:-r1x}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -q 2>&1 | tail -2 > $O/${TAG}_pytest_gpu.log
python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err
python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err
B="python bench.py --pairs 128 --steps 1 --warmup 1 --no-cpu-baseline --map-rows 0 --latency-iters 0 --stereo-pairs 16 --search-frames 16 --c4-frames 0"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_ncu_launches.csv $B > $O/ncu_launches.log 2>&1
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_fmaheavy.sum,sm__inst_executed_pipe_lsu.sum,sm__inst_executed_pipe_xu.sum --clock-control none --launch-skip 15 --launch-count 15 --csv --log-file $O/${TAG}_ncu_pipes.csv $B > $O/ncu_pipes.log 2>&1
ncu --set full --import-source on --clock-control none --launch-skip 15 --launch-count 15 -o $O/${TAG}_full -f $B > $O/ncu_full.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:stereo --launch-skip 2 --launch-count 2 -o $O/${TAG}_stereo -f $B > $O/ncu_stereo.log 2>&1
ncu -i $O/${TAG}_full.ncu-rep --page raw --csv > $O/${TAG}_full_raw.csv 2>/dev/null
ncu -i $O/${TAG}_stereo.ncu-rep --page raw --csv > $O/${TAG}_stereo_raw.csv 2>/dev/null
python tools/kernel_traffic.py $O/${TAG}_full_raw.csv 256 $TAG $O/${TAG}_ncu_pipes.csv > $O/${TAG}_kernel_traffic.json 2> $O/kernel_traffic.err
cuobjdump -sass orbslam3lib_b200/liborbfe.so | grep -E "Function :|UTCIMMA|LDTM|UTCBAR|UTMALDG|UBLKCP" | awk '/Function :/{f=$0; next} {c[f" | "$2]++} END{for(k in c) print c[k], k}' | sort -k2 > $O/${TAG}_sass_blackwell.txt
tail -2 $O/${TAG}_pytest_gpu.log; tail -c 600 $O/${TAG}_bench.json; ls -la $O | head -40
