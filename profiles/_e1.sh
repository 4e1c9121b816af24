N=5 ENVS=4096 TAG=r2_k_step_n5 timeout 900 bash profiles/r2_ncu.sh 2>&1 | tail -1
