# ncu durations of the six tower GEMMs of one Wide&Deep step: bash profiles/lin_ncu.sh [ENV=VAL ...]
for kv in "$@"; do export "$kv"; done
python profiles/run_wd_step.py > gpurun_out/p_wd.log 2>&1 ; ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -k regex:linear_tcgen05 --csv --log-file gpurun_out/lin_x.csv python profiles/run_wd_step.py > gpurun_out/n_wd.log 2>&1
(grep "gpu__time_duration" gpurun_out/lin_x.csv | awk -F'","' '{printf "%s ", $NF}' | tr -d '"'; echo " <- $*") >> gpurun_out/lin_all.txt
