"""Multi-GPU check of the sharded long-form Griffin-Lim (run under torchrun, one process per GPU):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/longform_gl_mp.py [seconds] [iters]
Every rank runs its fragment with seams linked over NVLink peer mappings; rank 0 gathers, compares with its own single-GPU
run of the whole spectrogram (SNR) and times both.  Writes gpurun_out/longform_gl_mp_<world>.json."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from unast_b200 import dist as D, longform as LF, synth, utils as U

def snr(a, b):
    e = a.astype(np.float64) - b
    return float(10 * np.log10((a.astype(np.float64) ** 2).sum() / max((e ** 2).sum(), 1e-300)))

secs = int(sys.argv[1]) if len(sys.argv) > 1 else 60
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 30
rank, local_rank, world = D.env_world()
torch.cuda.set_device(local_rank)
D.init_from_env("nccl")
eng = U.get_engine()
w = np.concatenate([synth.speech_like(10.0, 80 + i) for i in range((secs + 9) // 10)])[:secs * 22050]
(mel, mag), = U.spectrograms_from_waves([w], do_trim=False)
T = mag.shape[0]
sh = LF.ShardedGriffinLim(T, rank, world, canonical=False)        # the speed-optimised cut
rows = torch.from_numpy(np.ascontiguousarray(mag[sh.a:sh.b])).cuda()
raw = sh.run(rows, iters)
eng.check()
times = []
for _ in range(3):
    D.barrier(); torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record(); raw = sh.run(rows, iters); t1.record(); torch.cuda.synchronize()
    times.append(D.max_over_ranks(t0.elapsed_time(t1)))
eng.check()
wav, se = sh.gather_deemph_trim(raw)
s, e = se if se is not None else (0, 0)
out = None
if rank == 0:
    mag_d = torch.from_numpy(mag).cuda()
    single = LF.griffinlim_fragments_emulated(mag_d, 1, iters)
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record(); single = LF.griffinlim_fragments_emulated(mag_d, 1, iters); t1.record(); torch.cuda.synchronize()
    L = 275 * (T - 1)
    ref_wav, b = eng.deemph_trim(single, torch.tensor([0, L], dtype=torch.int64, device="cuda"), L)
    rs, re = (int(v) for v in b.cpu().numpy()[0])
    out = {"world": world, "seconds": secs, "frames": int(T), "iters": iters, "sharded_ms": min(times), "single_gpu_ms": t0.elapsed_time(t1),
           "trim_equal": bool((s, e) == (rs, re)), "snr_db_vs_single_gpu": snr(ref_wav.cpu().numpy(), wav.cpu().numpy()),
           "per_iteration_us": 1e3 * min(times) / (iters + 1)}
    print(json.dumps(out))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "longform_gl_mp_%d.json" % world), "w"), indent=1)
sh.close()
D.barrier()
if world > 1:
    torch.distributed.destroy_process_group()
