import sys, torch
sys.path.insert(0, "tests")
import support as S
from sparse_conv_gpis_tungsten_b200 import render as R
sc = S.appendix_a_scene(S.c4_scene_medium(), 1024, 128, quad=True)
stream = torch.cuda.current_stream()
for world in (1, 2, 4, 8):
    for rank in (0, world - 1):
        wr = R.WavefrontRenderer(sc, device=0, rank=rank, world=world, sharding="interleaved"); wr.render(1); wr.close()
        wr = R.WavefrontRenderer(sc, device=0, rank=rank, world=world, sharding="interleaved"); wr.prepare(128); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream); wr.render(128); b.record(stream); torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        chunk = max(1, min(128, wr.max_paths // wr.n_pixels))
        print("world %d rank %d: %.1f ms  x world = %.1f  (spp per chunk %d, %d chunks)" % (world, rank, ms, ms * world, chunk, 128 // chunk), flush=True)
        wr.close()
