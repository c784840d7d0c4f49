#!/bin/bash
# One gpurun call = a list of steps, each under its own timeout, each logging to gpurun_out/<tag>.log.
# A failing step does not stop the session (a hung or trapped kernel kills only its own process).
#   gpurun --timeout 1500 -- 'bash tools/gpu_session.sh probe tests bench'
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
step() {  # tag, timeout seconds, command...
    local tag=$1 to=$2; shift 2
    local t0=$(date +%s)
    timeout -k 10 "$to" "$@" > "gpurun_out/$tag.log" 2>&1
    local rc=$?
    echo "== $tag rc=$rc $(( $(date +%s) - t0 ))s"; tail -n 6 "gpurun_out/$tag.log" | cut -c1-400
}
for what in "$@"; do
case $what in
probe)
    step probe_round 300 python tools/gpu_probe_round.py ;;
sanitize_closed)
    PROBE_ONLY="n200" step sanitize_memcheck 600 compute-sanitizer --tool memcheck --print-limit 20 python tools/gpu_probe_round.py ;;
kernels)
    step t_kernels 600 python -m pytest tests/test_gpu_kernels.py -x -q ;;
tests)
    step t_kernels 600 python -m pytest tests/test_gpu_kernels.py -x -q
    step t_paths 900 python -m pytest tests/test_gpu_paths.py -x -q
    step t_glm 900 python -m pytest tests/test_gpu_glm_and_configs.py -x -q
    step t_wide 900 python -m pytest tests/test_gpu_wide_binomial.py -x -q
    step t_round2 900 python -m pytest tests/test_gpu_round2.py tests/test_zz_no_alloc.py -x -q ;;
alltests)
    step t_all 1500 python -m pytest tests -m gpu -x -q ;;
rbench)
    step round_bench 600 python tools/round_bench.py --cases 1x1,5x1,5x10 --lists 32,48,64,96,128 --skips 0 ;;
ncu_rb)
    step ncu_rb 600 ncu --set full --clock-control none --import-source on -k regex:irls_round_kernel -s 2 -c 1 -o gpurun_out/ncu_rb python tools/round_bench.py --cases 5x10 --lists 32 --skips 0 --iters 3
    ncu -i gpurun_out/ncu_rb.ncu-rep --page source --csv > gpurun_out/ncu_rb_source.csv 2>/dev/null
    ncu -i gpurun_out/ncu_rb.ncu-rep --page raw --csv > gpurun_out/ncu_rb_raw.csv 2>/dev/null
    rm -f gpurun_out/ncu_rb.ncu-rep ;;
trace)
    for c in "1x1 64" "5x10 32" "5x10 64"; do set -- $c; tag=$(echo $1_$2)
      ATE_B200_LIB=$PWD/ate_replication_causalml_b200/libate_b200_trace.so ATE_ROUND_TRACE_OUT=gpurun_out/trace_$tag.csv step trace_$tag 300 python tools/round_bench.py --cases $1 --lists $2 --skips 0 --iters 2
      python tools/round_trace.py gpurun_out/trace_$tag.csv 60 76
    done ;;
hbm)
    step hbm_passes 600 python tools/hbm_passes.py ;;
ncu_hbm)
    step ncu_hbm 900 ncu --set full --clock-control none -k regex:'col_moments_kernel|col_apply|build_design|expand_col|aipw_score|bootstrap_kernel|gather_rows|cvdev|predict_rows|rk_rank|rk_hist|scatter_kernel|inner_key' -c 40 -o gpurun_out/ncu_hbm python tools/hbm_passes.py
    ncu -i gpurun_out/ncu_hbm.ncu-rep --page raw --csv > gpurun_out/ncu_hbm_raw.csv 2>/dev/null
    rm -f gpurun_out/ncu_hbm.ncu-rep ;;
sass)
    python tools/sass_summary.py > gpurun_out/sass_summary.txt 2>&1 ;;
smoke)
    step smoke 300 python -c "import __graft_entry__ as g; g.smoke()" ;;
bench)
    step bench_c1 900 python bench.py --steps 2 --warmup 3
    grep -h '^{' gpurun_out/bench_c1.log | tail -1 > gpurun_out/bench_c1.json ;;
benchq)
    ATE_PROFILE_CSV=gpurun_out/round_launches.csv ATE_ROUND_DEBUG=1 step bench_c1q 600 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e
    grep -h '^{' gpurun_out/bench_c1q.log | tail -1 > gpurun_out/bench_c1q.json ;;
ref)
    step bench_ref 600 python bench.py --impl reference --steps 1 --warmup 1 --linearity
    grep -h '^{' gpurun_out/bench_ref.log | tail -1 > gpurun_out/bench_ref.json ;;
c2)
    step bench_c2 600 python bench.py --workload C2 --steps 3 --warmup 3
    grep -h '^{' gpurun_out/bench_c2.log | tail -1 > gpurun_out/bench_c2.json ;;
c3)
    step bench_c3 900 python bench.py --workload C3 --steps 2 --warmup 1
    grep -h '^{' gpurun_out/bench_c3.log | tail -1 > gpurun_out/bench_c3.json ;;
c4)
    step bench_c4 1500 python bench.py --workload C4 --steps 1 --warmup 1 --no-e2e
    grep -h '^{' gpurun_out/bench_c4.log | tail -1 > gpurun_out/bench_c4.json ;;
ncu_list)
    step ncu_launches 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/ncu_launches.csv python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e ;;
mgcheck)
    NG=${NGPU:-2}
    step mg${NG}_check 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29511 tools/multi_gpu_check.py ;;
mgbench)
    NG=${NGPU:-2}
    step bench_c1_n$NG 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $NG --steps 2 --warmup 3
    grep -h '^{' gpurun_out/bench_c1_n$NG.log | tail -1 > gpurun_out/bench_c1_n$NG.json ;;
mgref)
    NG=${NGPU:-2}
    step bench_ref_n$NG 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus $NG --steps 1 --warmup 1 ;;
mgc4)
    NG=${NGPU:-2}
    step bench_c4_n$NG 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $NG --workload C4 --steps 1 --warmup 1 --no-e2e
    grep -h '^{' gpurun_out/bench_c4_n$NG.log | tail -1 > gpurun_out/bench_c4_n$NG.json ;;
ncu_round)
    step ncu_round 900 ncu --set full --clock-control none --import-source on -k regex:irls_round_kernel -s 150 -c 3 -o gpurun_out/ncu_round python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e
    ncu -i gpurun_out/ncu_round.ncu-rep --page raw --csv > gpurun_out/ncu_round_raw.csv 2>/dev/null
    ncu -i gpurun_out/ncu_round.ncu-rep --page source --csv > gpurun_out/ncu_round_source.csv 2>/dev/null
    rm -f gpurun_out/ncu_round.ncu-rep ;;
ncu_wide)
    step ncu_wide 600 ncu --set full --clock-control none --import-source on -k regex:irls_round_kernel -s 2 -c 1 -o gpurun_out/ncu_wide python tools/round_bench.py --cases 5x10 --lists ${WIDE_LIST:-128} --skips 0 --iters 3
    ncu -i gpurun_out/ncu_wide.ncu-rep --page source --csv > gpurun_out/ncu_wide_source.csv 2>/dev/null
    ncu -i gpurun_out/ncu_wide.ncu-rep --page raw --csv > gpurun_out/ncu_wide_raw.csv 2>/dev/null
    rm -f gpurun_out/ncu_wide.ncu-rep ;;
*)
    echo "unknown step $what" ;;
esac
done
# gpurun merges at most 64 MiB back: drop the largest files until the directory fits
while [ "$(du -sm gpurun_out | cut -f1)" -gt 56 ]; do
    big=$(ls -S gpurun_out | head -1); echo "dropping gpurun_out/$big ($(du -sm gpurun_out/$big | cut -f1) MiB)"; rm -f "gpurun_out/$big"
done
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,temperature.gpu,power.draw --format=csv,noheader
