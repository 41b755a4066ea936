import os, sys, torch, numpy as np
sys.path.insert(0, '/root/repo')
from madmom_b200 import recipe as R
from madmom_b200.batch import BatchFrontEnd
from madmom_b200.synth import make_batch_torch
pcm = make_batch_torch(1024, 30.0, device='cuda', seed=1)
lengths = [pcm.shape[1]] * 1024
for rec in list(R.RNN_BEAT) + list(R.RNN_ONSET):
    fe = BatchFrontEnd([rec], np.int16)
    info = fe.plans[0].info
    units, T = fe.units(lengths); rows = int(sum(T))
    outs = ('log', 'diff')
    out = fe.run_device(pcm, lengths, outs, units=units, rows=rows)
    for _ in range(3): fe.run_device(pcm, lengths, outs, out=out, units=units, rows=rows)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(5): fe.run_device(pcm, lengths, outs, out=out, units=units, rows=rows)
    b.record(); torch.cuda.synchronize()
    print('N', rec.frame_size, 'bands', info['num_bands'], 'warps', info['warps'], 'smem', info['smem_bytes'], 'fpr', info['frames_per_round'], 'rows', rows, 'ms %.3f' % (a.elapsed_time(b) / 5), flush=True)
    del fe, out
