set -x
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum"
python profiles/run_step.py > gpurun_out/p_step.log 2>&1 && ncu --metrics $M --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches7.csv python profiles/run_step.py > gpurun_out/n_step.log 2>&1; echo rc=$?
python profiles/run_wd_step.py > gpurun_out/p_wd.log 2>&1 && ncu --metrics $M --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_wd5.csv python profiles/run_wd_step.py > gpurun_out/n_wd.log 2>&1; echo rc=$?
python profiles/run_wgrad3x3.py > gpurun_out/p_wg.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:wgrad3x3 -o gpurun_out/prof_wgrad3x3 -f python profiles/run_wgrad3x3.py > gpurun_out/n_wg.log 2>&1; echo rc=$?
python profiles/run_allreduce_emulated.py > gpurun_out/p_ar.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"allreduce|alltoall" -s 8 -c 4 -o gpurun_out/prof_collectives2 -f python profiles/run_allreduce_emulated.py > gpurun_out/n_ar.log 2>&1; echo rc=$?
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"emb_|linear_tcgen05" -c 8 -o gpurun_out/prof_wd2 -f python profiles/run_wd_step.py > gpurun_out/n_wd2.log 2>&1; echo rc=$?
python profiles/gemm_bench.py > gpurun_out/gemm_bench.log 2>&1; tail -5 gpurun_out/gemm_bench.log
tail -2 gpurun_out/p_ar.log gpurun_out/n_ar.log
