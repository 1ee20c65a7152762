O=gpurun_out/d9; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512"
timeout 400 $TR bench.py --gpus 8 --steps 20 --warmup 5 > $O/bench_cfg3_n8.json 2> $O/bench_cfg3_n8.err; tail -c 900 $O/bench_cfg3_n8.json; tail -2 $O/bench_cfg3_n8.err
timeout 400 $TR bench.py --gpus 8 --steps 10 --warmup 3 --workload aishell_train_step_B32_T500_U50 > $O/bench_train_n8.json 2> $O/bench_train_n8.err; tail -c 900 $O/bench_train_n8.json; tail -2 $O/bench_train_n8.err
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_cfg3_n1.json 2> $O/bench_cfg3_n1.err; tail -c 300 $O/bench_cfg3_n1.json
