N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29511 bench.py --gpus $N > gpurun_out/bench_r02_n${N}.json 2> gpurun_out/bench_r02_n${N}.err; echo "bench rc $?"
timeout 300 $TR --master-port 29512 tools/bench_configs.py mc --steps 20 > gpurun_out/bench_r02_cfg3_mc_n${N}.json 2> gpurun_out/cfg3_n${N}.err; echo "mc rc $?"; tail -1 gpurun_out/cfg3_n${N}.err
timeout 400 $TR --master-port 29513 tools/bench_configs.py sweep > gpurun_out/cfg5_sweep_r02_n${N}.jsonl 2> gpurun_out/sweep_n${N}.err; echo "sweep rc $?"
timeout 400 $TR --master-port 29514 tools/bench_configs.py sweep_mc --max-restarts 1024 --steps 20 > gpurun_out/cfg5_sweep_mc_r02_n${N}.jsonl 2> gpurun_out/sweep_mc_n${N}.err; echo "sweep_mc rc $?"
python - <<PY
import json
for f in ["gpurun_out/bench_r02_n${N}.json","gpurun_out/bench_r02_cfg3_mc_n${N}.json"]:
    for l in open(f):
        if l.startswith("{"):
            j=json.loads(l); print(f, j.get("value"), j.get("ms_per_step"), (j.get("parity_selfcheck") or {}).get("ok"), (j.get("roofline") or {}).get("frac"))
print(sum(1 for _ in open("gpurun_out/cfg5_sweep_r02_n${N}.jsonl")), sum(1 for _ in open("gpurun_out/cfg5_sweep_mc_r02_n${N}.jsonl")))
PY
