#!/bin/bash
# launch lists (durations) for the three step kinds + one full ncu capture of the hot kernels of each
mkdir -p gpurun_out
for cfg in mnist mnist_chain wide sparse wd; do
  python tools/profile_step.py --config $cfg --steps 4 > gpurun_out/plain_$cfg.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$cfg.csv python tools/profile_step.py --config $cfg --steps 4 > gpurun_out/ncu_l_$cfg.log 2>&1
  echo "launch list $cfg: $?"
done
python tools/profile_step.py --config wide --steps 2 > gpurun_out/plain_wide2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'gemm_tcgen05' -s 30 -c 4 -o gpurun_out/prof_wide -f python tools/profile_step.py --config wide --steps 2 > gpurun_out/ncu_wide.log 2>&1
echo "ncu wide: $?"
python tools/profile_step.py --config mnist --steps 6 > gpurun_out/plain_mnist2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'mlp_fused' -s 3 -c 1 -o gpurun_out/prof_mnist -f python tools/profile_step.py --config mnist --steps 6 > gpurun_out/ncu_mnist.log 2>&1
echo "ncu mnist: $?"
python tools/profile_step.py --config sparse --steps 6 > gpurun_out/plain_sparse2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'sparse_lr' -s 2 -c 2 -o gpurun_out/prof_sparse -f python tools/profile_step.py --config sparse --steps 6 > gpurun_out/ncu_sparse.log 2>&1
echo "ncu sparse: $?"
python tools/profile_step.py --config mnist_chain --steps 6 > gpurun_out/plain_mnist3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'mlp_tail|push_sync|gemm_tcgen05' -s 12 -c 4 -o gpurun_out/prof_mnist_chain -f python tools/profile_step.py --config mnist_chain --steps 6 > gpurun_out/ncu_mnist_chain.log 2>&1
echo "ncu mnist chain: $?"
python tools/profile_step.py --config wd --steps 6 > gpurun_out/plain_wd2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'sparse_wd' -s 2 -c 2 -o gpurun_out/prof_wd -f python tools/profile_step.py --config wd --steps 6 > gpurun_out/ncu_wd.log 2>&1
echo "ncu wd: $?"
ls -la gpurun_out/*.ncu-rep
python tools/prof_wd_owner.py > gpurun_out/plain_wd_owner.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'wd_' -s 15 -c 5 -o gpurun_out/prof_wd_owner -f python tools/prof_wd_owner.py > gpurun_out/ncu_wd_owner.log 2>&1
echo "ncu wd owner: $?"
python tools/prof_cluster.py > gpurun_out/plain_cluster.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'mlp_cluster' -s 2 -c 1 -o gpurun_out/prof_cluster -f python tools/prof_cluster.py > gpurun_out/ncu_cluster.log 2>&1
echo "ncu cluster: $?"
