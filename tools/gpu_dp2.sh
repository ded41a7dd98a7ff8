mkdir -p gpurun_out
export NCCL_DEBUG=WARN
timeout 200 python -m pytest tests/test_gpu_dp.py -x -q 2>&1 | tail -3
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29712 bench.py --gpus 2 > gpurun_out/bench_dp2_graph.json 2> gpurun_out/bench_dp2_graph.err; echo graph rc=$?; cut -c1-330 gpurun_out/bench_dp2_graph.json; tail -5 gpurun_out/bench_dp2_graph.err
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 --no-graph > gpurun_out/bench_dp2_eager.json 2> gpurun_out/bench_dp2_eager.err; echo eager rc=$?; cut -c1-330 gpurun_out/bench_dp2_eager.json
