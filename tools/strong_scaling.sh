set -x
for n in 2 4 8; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --gpus $n --workload rans --nx 4001 --ny 4001 --strong --steps 5 --warmup 3 --no-cpu-baseline 2> gpurun_out/strong32m_n$n.err | tail -1 > gpurun_out/strong32m_n$n.json
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29650 bench.py --gpus 8 --workload rans --nx 1415 --ny 1415 --strong --partition graph --steps 5 --warmup 3 --no-cpu-baseline 2> gpurun_out/graph4m_n8.err | tail -1 > gpurun_out/graph4m_n8.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29651 bench.py --gpus 8 --workload rans --nx 1415 --ny 1415 --strong --partition strips --steps 5 --warmup 3 --no-cpu-baseline 2> gpurun_out/strips4m_n8.err | tail -1 > gpurun_out/strips4m_n8.json
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/strong32m_n*.json"))+["gpurun_out/graph4m_n8.json","gpurun_out/strips4m_n8.json"]:
    try:
        d=json.load(open(f)); print(f, d["n_gpus"], round(d["ms_per_step"],3), "%.4g"%d["value"], round(d["e2e"]["ms_per_step"],2), d["config"]["points_per_gpu"], d["gpu_launches"])
    except Exception as e: print(f, "ERR", e)
PY
