T=r3g; O=gpurun_out
timeout 300 python tools/profile_case.py 30 0 > $O/${T}_plain.log 2>&1 && \
PROFILE_REPS=1 timeout 900 ncu --set full --clock-control none --import-source on --replay-mode application -k regex:avr_job_kernel \
    -o /tmp/${T}_full -f python tools/profile_case.py 30 0 > $O/${T}_ncu.log 2>&1
ncu -i /tmp/${T}_full.ncu-rep --page source --print-source cuda --csv 2> $O/${T}_src.err | gzip -9 > $O/${T}_source_cuda.csv.gz
ncu -i /tmp/${T}_full.ncu-rep --page source --csv 2>> $O/${T}_src.err | gzip -9 > $O/${T}_source_sass.csv.gz
ls -la $O | tail -5; tail -2 $O/${T}_ncu.log
