# usage: tools/scale_r2.sh <tag>   (under gpurun --gpus 8) -- the default workload at N = 1, 2, 4, 8 back to back, as the driver runs it
T=$1; O=gpurun_out
python bench.py --gpus 1 --steps 2 --warmup 1 --no-cpu-baseline > $O/scale_n1_$T.json 2> $O/scale_n1_$T.err
for N in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + N)) bench.py --gpus $N --steps 2 --warmup 1 \
      > $O/scale_n${N}_$T.json 2> $O/scale_n${N}_$T.err || tail -5 $O/scale_n${N}_$T.err
done
python - <<'PY'
import json, glob, os
t = os.environ.get("T", "")
base = None
for n in (1, 2, 4, 8):
    p = [f for f in glob.glob(f"gpurun_out/scale_n{n}_*.json")]
    if not p: continue
    try: d = json.load(open(sorted(p)[-1]))
    except Exception as e: print(n, "no result", e); continue
    base = base or d["ms_per_step"]
    print(f"N={n}: {d['ms_per_step']:.1f} ms, {d['value']:.2f} Msamples/s, efficiency {base / d['ms_per_step'] / n:.3f}, checksum {d['checksum']} / {d['checksum_sigma2']}, "
          f"parity {d['parity']['pass'] if d['parity'] else None} / {d['parity_sigma2']['pass'] if d['parity_sigma2'] else None} straddles {d['parity']['straddles_band_boundary'] if d['parity'] else None}, "
          f"rank step ms {d['ranks']['step_ms']['min']:.1f}..{d['ranks']['step_ms']['max']:.1f}, e2e {d['e2e']['value']:.2f}")
PY
