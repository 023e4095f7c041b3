#!/bin/bash
# A/B of the global (out-of-shared-memory) path of k_render_pipe: new libcrt.so against libcrt_ab_head.so, plus the GPU tests that
# exercise trees that change on the device (LBVH build, refit, edits) and the full-size scenes. $1 = tag.
T=${1:-r02big}
mkdir -p gpurun_out
out=gpurun_out/${T}_ab.txt; : > $out
A=/root/repo/cudaraytracer_b200/libcrt_ab_head.so; B=/root/repo/cudaraytracer_b200/libcrt.so
qp() { timeout 200 python tools/quick_perf.py "$@" 2>&1 | grep -E "Mrays|sha1" | tail -2; }
echo "== sanity stress 640x360 4 spp depth 100 (sha1 must agree)" >> $out
for lib in $A $B; do CRT_LIB_PATH=$lib qp stress 640 360 4 100 3 1 >> $out; done
if ! tail -1 $out | grep -q sha1; then echo "NEW BUILD FAILED THE SANITY RUN" >> $out; cat $out; exit 1; fi
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_lbvh.py tests/test_gpu_edit.py tests/test_gpu_fullsize.py -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_pytest.log
tail -4 gpurun_out/${T}_pytest.log >> $out
for rep in 1 2; do for lib in $A $B; do
  echo "== $(basename $lib) grid 4K 4 spp" >> $out; CRT_LIB_PATH=$lib qp grid 3840 2160 4 50 3 2 >> $out
  echo "== $(basename $lib) stress 4K 4 spp depth 100" >> $out; CRT_LIB_PATH=$lib qp stress 3840 2160 4 100 3 2 >> $out
done; done
for lib in $A $B; do echo "== $(basename $lib) stress 8K 2 spp depth 100" >> $out; CRT_LIB_PATH=$lib qp stress 7680 4320 2 100 3 2 >> $out; done
for kb in 8 48; do echo "== new CRT_PIPE_PREFIX_KB=$kb grid" >> $out; CRT_PIPE_PREFIX_KB=$kb CRT_LIB_PATH=$B qp grid 3840 2160 4 50 3 2 >> $out; done
for t in 4 8; do echo "== new CRT_PIPE_THRESH=$t grid" >> $out; CRT_PIPE_THRESH=$t CRT_LIB_PATH=$B qp grid 3840 2160 4 50 3 2 >> $out; done
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum
for lib in $A $B; do
CRT_LIB_PATH=$lib timeout 600 ncu --metrics $M --clock-control none -k regex:k_render_pipe -s 1 -c 1 --csv --log-file gpurun_out/${T}_metrics_$(basename $lib .so).csv python tools/quick_perf.py grid 3840 2160 4 50 3 1 > /dev/null 2>&1
done
grep -v "^accum" $out; grep "^accum" $out | sort | uniq -c
