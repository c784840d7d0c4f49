for T in 10 8; do
ATE_PLAN_MINT=$T ATE_STEP_WATCHDOG=1 timeout 100 python bench.py --steps 2 --warmup 3 > gpurun_out/v5_T$T.json 2> gpurun_out/v5_T$T.err; echo exit=$? >> gpurun_out/v5_T$T.json
done
