set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 1500 python -m pytest tests/test_distributed.py -m gpu -x -q > gpurun_out/r2_t8.log 2>&1; tail -4 gpurun_out/r2_t8.log
for mc in 1 0; do
  python tools/nvlink_delta.py 0 > gpurun_out/nvl_before_$mc.txt
  B200GRU_TP_MC=$mc timeout 600 $TR --master-port 2970$mc bench.py --gpus 8 --parallel tp --steps 20 > gpurun_out/r2_tp8_mc$mc.json 2> gpurun_out/r2_tp8_mc$mc.err
  python tools/nvlink_delta.py 0 > gpurun_out/nvl_after_$mc.txt
  tail -c 1500 gpurun_out/r2_tp8_mc$mc.json
done
timeout 600 $TR --master-port 29711 bench.py --gpus 8 --exact-names 0 > gpurun_out/r2_dp8.json 2> gpurun_out/r2_dp8.err; tail -c 900 gpurun_out/r2_dp8.json
timeout 600 $TR --master-port 29712 benchmarks/sweep.py --sizes 1024,4096,16384,65536,262144 --steps 10 2>&1 | grep n_names | cut -c1-150 | tee gpurun_out/r2_sweep8.log
