# ncu --set full capture of the step kernel (one combined launch), after the same command ran clean without ncu
mkdir -p gpurun_out
N=${N:-5}; ENVS=${ENVS:-4096}; TAG=${TAG:-r2_k_step_n5}
python profiles/profile_step.py $ENVS $N 16 > gpurun_out/p_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_step --launch-skip 26 --launch-count 2 -o gpurun_out/$TAG -f python profiles/profile_step.py $ENVS $N 16 > gpurun_out/ncu_$TAG.log 2>&1
cat gpurun_out/p_$TAG.log; tail -3 gpurun_out/ncu_$TAG.log; ls -la gpurun_out/$TAG.ncu-rep
