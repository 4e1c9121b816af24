N=8 ENVS=16384 TAG=r2_k_step_n8 timeout 1200 bash profiles/r2_ncu.sh 2>&1 | tail -1
