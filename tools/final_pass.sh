#!/bin/bash
# The round's evidence in one GPU call (run under gpurun from the repo root): bench lines of every workload, the ncu launch
# list of the default bench command, one ncu --set full capture per workload's march kernel.  Outputs under gpurun_out/;
# summarise here with tools/ncu_summary.py and tools/update_traffic.py.
set -u
R=${1:-r02}
O=gpurun_out
: > $O/${R}_bench_final.jsonl
for w in c2 c1 c3 c4s128 c4 c4s1024 c5 bake; do
  python bench.py --workload $w --steps 20 --warmup 5 --no-extra 2>>$O/${R}_bench_final.err | tail -1 >> $O/${R}_bench_final.jsonl
done
python bench.py --impl reference --steps 2 --warmup 1 2>>$O/${R}_bench_final.err | tail -1 >> $O/${R}_bench_final.jsonl
python bench.py --impl reference --workload c1 --steps 2 --warmup 1 2>>$O/${R}_bench_final.err | tail -1 >> $O/${R}_bench_final.jsonl
python bench.py --steps 30 --warmup 5 > $O/${R}_bench_default_pre.json 2>>$O/${R}_bench_final.err
# launch list of the default command (after it has exited 0 without ncu, above)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${R}_launches_default.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $O/ncu_launches.log 2>&1
cap() {  # workload, kernel regex, skip: capture, summarise and record the DRAM traffic here (the reports are too large to travel)
  ncu --set full --clock-control none -k regex:$2 -s $3 -c 1 -o $O/${R}_full_$1 python bench.py --workload $1 --steps 2 --warmup 3 --no-cpu-baseline --no-extra > $O/ncu_$1.log 2>&1
  python tools/ncu_summary.py $O/${R}_full_$1.ncu-rep > $O/${R}_final_$1.txt
  python tools/update_traffic.py $1 $O/${R}_full_$1.ncu-rep $2 >> $O/traffic.log 2>&1
  rm -f $O/${R}_full_$1.ncu-rep
}
cap c2 march_kernel 4
cap c3 march_kernel 4
cap c1 march_tex_split 4
cap c4s128 march_kernel 4
cap c5 march_kernel 4
cp profiles/traffic.json $O/traffic.json
# the default bench line again, now that traffic.json belongs to this kernel source (roofline.traffic filled in)
python bench.py --steps 30 --warmup 5 > $O/${R}_bench_default.json 2>>$O/${R}_bench_final.err
ls -la $O | tail -20
