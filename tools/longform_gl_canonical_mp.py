"""Multi-GPU check (torchrun, one process per GPU): canonical sharded Griffin-Lim against the canonical single-GPU run, bit for bit,
at several iteration counts and twice in a row (short calls expose a missing hand-over at the linked seams).  Args: [seconds] [chunk]."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from unast_b200 import dist as D, longform as LF, synth, utils as U
rank, local_rank, world = D.env_world()
torch.cuda.set_device(local_rank)
D.init_from_env("nccl")
eng = U.get_engine()
secs = int(sys.argv[1]) if len(sys.argv) > 1 else 60
w = np.concatenate([synth.speech_like(10.0, 80 + i) for i in range((secs + 9) // 10)])[:secs * 22050]
(mel, mag), = U.spectrograms_from_waves([w], do_trim=False)
T = mag.shape[0]
sh = LF.ShardedGriffinLim(T, rank, world, canonical=True)
rows = torch.from_numpy(np.ascontiguousarray(mag[sh.a:sh.b])).cuda()
if len(sys.argv) > 2:                                       # rows recomputed from this rank's chunk of the waveform, as bench.py does
    from unast_b200.engine import HOP
    pcm = synth.to_pcm16(w)
    (mel_i, mag_i), = U.spectrograms_from_waves([pcm], do_trim=False)
    n = pcm.size
    k0, k1, kk = LF.chunk_for_range(n, sh.a, sh.b)
    ch = torch.from_numpy(pcm[k0:k1].copy()).cuda()
    o2 = torch.tensor([0, k1 - k0], dtype=torch.int64, device="cuda")
    b2 = torch.tensor([[0, k1 - k0]], dtype=torch.int32, device="cuda")
    f2 = eng.frame_offsets(b2)
    _, mg2 = eng.frontend(ch, o2, b2, f2, int(f2.cpu().numpy()[-1]), precision="f64")
    rows = mg2[kk].contiguous()
    mag = mag_i
    print("rank", rank, "rows equal to the slice of the whole-waveform run:", bool(np.array_equal(rows.cpu().numpy(), mag[sh.a:sh.b])), flush=True)
mag_d = torch.from_numpy(mag).cuda()
for iters in ((0, 1, 2, 10, 100) if secs <= 60 else (0, 2, 100)):
    outs = []
    for rep in range(2):
        raw = sh.run(rows, iters)
        full = sh.gather_rows(raw) if world > 1 else raw
        if rank == 0:
            outs.append(full.cpu().numpy())
    if rank == 0:
        single = LF.griffinlim_fragments_emulated(mag_d, 1, iters, canonical=True).cpu().numpy()
        d = outs[0] != single
        e = outs[0].astype(np.float64) - single
        print("iters", iters, "range", (sh.a, sh.b), "run-to-run equal", bool(np.array_equal(outs[0], outs[1])), "mismatches", int(d.sum()),
              "first/last", (int(np.argmax(d)) if d.any() else -1, int(len(d) - 1 - np.argmax(d[::-1])) if d.any() else -1),
              "snr %.1f" % (10 * np.log10((single.astype(np.float64) ** 2).sum() / max((e ** 2).sum(), 1e-300))), "seam sample", 275 * sh.b, flush=True)
sh.close()
D.barrier()
torch.distributed.destroy_process_group()
