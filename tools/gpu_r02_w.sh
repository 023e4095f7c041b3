#!/bin/bash
# Round-2 visit w: the bank-group queues of k_render_pipe (libcrt.so) against the committed build (libcrt_ab_head.so), in ONE
# call (boxes differ between calls). Every command runs under its own timeout: a queue bug hangs, it does not crash.
T=${1:-r02w}
mkdir -p gpurun_out
out=gpurun_out/${T}_ab.txt; : > $out
A=/root/repo/cudaraytracer_b200/libcrt_ab_head.so; B=/root/repo/cudaraytracer_b200/libcrt.so
qp() { timeout 120 python tools/quick_perf.py "$@" 2>&1 | grep -E "Mrays|sha1" | tail -2; }
# 0. does the new build finish at all, and with the committed build's bits? (small frame first)
echo "== sanity 400x225 16 spp (sha1 must agree)" >> $out
for lib in $A $B; do CRT_LIB_PATH=$lib qp rtiow 400 225 16 50 3 1 >> $out; done
if ! tail -1 $out | grep -q sha1; then echo "NEW BUILD FAILED THE SANITY RUN" >> $out; cat $out; exit 1; fi
# 1. bit-identity + small pools
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/${T}_pytest_parity.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_pytest_parity.log
tail -3 gpurun_out/${T}_pytest_parity.log >> $out
# 2. A/B on the three scene classes
for rep in 1 2; do for lib in $A $B; do
  echo "== $(basename $lib) rtiow 1080p 64 spp" >> $out; CRT_LIB_PATH=$lib qp rtiow 1920 1080 64 50 3 3 >> $out
done; done
for lib in $A $B; do
  echo "== $(basename $lib) grid 4K 4 spp" >> $out; CRT_LIB_PATH=$lib qp grid 3840 2160 4 50 3 2 >> $out
  echo "== $(basename $lib) stress 4K 4 spp depth 100" >> $out; CRT_LIB_PATH=$lib qp stress 3840 2160 4 100 3 2 >> $out
  echo "== $(basename $lib) textured 1080p 64 spp (pipeline)" >> $out; CRT_LIB_PATH=$lib qp textured 1920 1080 64 50 3 2 >> $out
done
# 3. when is a class queue "full"? (new build only)
for f in 20 24 28 30 32; do echo "== new CRT_PIPE_FULL=$f rtiow" >> $out; CRT_PIPE_FULL=$f CRT_LIB_PATH=$B qp rtiow 1920 1080 64 50 3 2 >> $out; done
for t in 10 12 16 18; do echo "== new CRT_PIPE_THRESH=$t rtiow" >> $out; CRT_PIPE_THRESH=$t CRT_LIB_PATH=$B qp rtiow 1920 1080 64 50 3 2 >> $out; done
# 4. what a 1024-slot pool costs the committed build (would a shared-memory traversal stack fit?)
echo "== head CRT_PIPE_SLOTS=1024 rtiow" >> $out; CRT_PIPE_SLOTS=1024 CRT_LIB_PATH=$A qp rtiow 1920 1080 64 50 3 2 >> $out
echo "== new CRT_PIPE_SLOTS=1024 rtiow" >> $out; CRT_PIPE_SLOTS=1024 CRT_LIB_PATH=$B qp rtiow 1920 1080 64 50 3 2 >> $out
# 5. shared-memory wavefronts / conflicts / local-memory share of both builds (ncu, metrics only; never a bench number)
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__data_pipe_lsu_wavefronts.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared_op_atom.sum,l1tex__t_requests_pipe_lsu_mem_local_op_ld.sum,l1tex__t_requests_pipe_lsu_mem_local_op_st.sum,l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum,l1tex__t_sectors_pipe_lsu_mem_local_op_ld_lookup_hit.sum,l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum,l1tex__data_pipe_lsu_wavefronts_mem_lg.sum,smsp__inst_executed_op_shared_atom.sum,smsp__inst_executed_op_local_ld.sum,smsp__inst_executed_op_local_st.sum
for lib in $A $B; do
  CRT_LIB_PATH=$lib timeout 600 ncu --metrics $M --clock-control none -k regex:k_render_pipe -s 1 -c 1 --csv --log-file gpurun_out/${T}_metrics_$(basename $lib .so).csv python tools/quick_perf.py rtiow 1920 1080 64 50 3 1 > /dev/null 2>&1
done
cat $out
