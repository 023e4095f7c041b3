#!/bin/bash
# Round-2 visit y: A/B of a new libcrt.so against the committed build (libcrt_ab_head.so) in ONE call. $1 = tag.
T=${1:-r02y}
mkdir -p gpurun_out
out=gpurun_out/${T}_ab.txt; : > $out
A=/root/repo/cudaraytracer_b200/libcrt_ab_head.so; B=/root/repo/cudaraytracer_b200/libcrt.so
qp() { timeout 120 python tools/quick_perf.py "$@" 2>&1 | grep -E "Mrays|sha1" | tail -2; }
echo "== sanity 400x225 16 spp (sha1 must agree)" >> $out
for lib in $A $B; do CRT_LIB_PATH=$lib qp rtiow 400 225 16 50 3 1 >> $out; done
if ! tail -1 $out | grep -q sha1; then echo "NEW BUILD FAILED THE SANITY RUN" >> $out; cat $out; exit 1; fi
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/${T}_pytest_parity.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_pytest_parity.log
tail -3 gpurun_out/${T}_pytest_parity.log >> $out
for rep in 1 2; do for lib in $A $B; do
  echo "== $(basename $lib) rtiow 1080p 64 spp" >> $out; CRT_LIB_PATH=$lib qp rtiow 1920 1080 64 50 3 3 >> $out
done; done
for lib in $A $B; do
  echo "== $(basename $lib) grid 4K 4 spp" >> $out; CRT_LIB_PATH=$lib qp grid 3840 2160 4 50 3 2 >> $out
  echo "== $(basename $lib) stress 4K 4 spp depth 100" >> $out; CRT_LIB_PATH=$lib qp stress 3840 2160 4 100 3 2 >> $out
  echo "== $(basename $lib) textured 1080p 64 spp (pipeline)" >> $out; CRT_LIB_PATH=$lib qp textured 1920 1080 64 50 3 2 >> $out
  echo "== $(basename $lib) default scene 1080p 64 spp (pipeline)" >> $out; CRT_LIB_PATH=$lib qp default 1920 1080 64 50 3 2 >> $out
done
for t in 10 12 16; do echo "== new CRT_PIPE_THRESH=$t rtiow" >> $out; CRT_PIPE_THRESH=$t CRT_LIB_PATH=$B qp rtiow 1920 1080 64 50 3 2 >> $out; done
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum
CRT_LIB_PATH=$B timeout 600 ncu --metrics $M --clock-control none -k regex:k_render_pipe -s 1 -c 1 --csv --log-file gpurun_out/${T}_metrics_new.csv python tools/quick_perf.py rtiow 1920 1080 64 50 3 1 > /dev/null 2>&1
grep -v "^accum" $out; grep "^accum" $out | sort | uniq -c
