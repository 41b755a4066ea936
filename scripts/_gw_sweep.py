import os, sys, torch, numpy as np
sys.path.insert(0, '/root/repo')
from madmom_b200 import recipe as R
from madmom_b200.batch import BatchFrontEnd
from madmom_b200.synth import make_batch_torch
pcm = make_batch_torch(1024, 30.0, device='cuda', seed=1)
lengths = [pcm.shape[1]] * 1024
for gw in (0,):
    for wmax in (16,):
        if wmax % gw: continue
        os.environ['MMB_GROUP_WARPS'] = str(gw); os.environ['MMB_WARPS'] = str(wmax)
        fe = BatchFrontEnd([R.LOGFILTSPECFLUX], np.int16)
        info = fe.plans[0].info
        units, T = fe.units(lengths); rows = int(sum(T))
        out = fe.run_device(pcm, lengths, ('diff', 'odf'), units=units, rows=rows)
        for _ in range(3): fe.run_device(pcm, lengths, ('diff', 'odf'), out=out, units=units, rows=rows)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); a.record()
        for _ in range(10): fe.run_device(pcm, lengths, ('diff', 'odf'), out=out, units=units, rows=rows)
        b.record(); torch.cuda.synchronize()
        print('gw', gw, 'wmax', wmax, 'warps', info['warps'], 'smem', info['smem_bytes'], 'fpr', info['frames_per_round'], 'chunk', info['chunk_frames'], 'ms %.3f' % (a.elapsed_time(b) / 10), flush=True)
        del fe, out
