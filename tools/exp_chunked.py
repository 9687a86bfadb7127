"""Experiment: G+D forward over the swarm in L2-sized particle chunks (activations stay in the 126 MB L2)
vs one pass over all 65,536 particles.  Chunk c runs on stream c % nbuf with its own activation buffers."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gan-discovery-pso_b200"))
import torch
import bench
from psob import engine

P = 65536
gen, disc = bench.make_models()
z = torch.zeros(P, 128, dtype=torch.bfloat16, device="cuda")
engine.pack_latent(torch.randn(P, 100, device="cuda"), z)
for chunk, nbuf in ((65536, 1), (16384, 2), (8192, 2), (4096, 2), (2048, 2), (2048, 3), (1024, 2), (1024, 4)):
    nets = []
    for _ in range(nbuf):
        g = engine.GeneratorNet(gen.state_dict(), chunk); d = engine.DiscriminatorNet(disc.state_dict(), chunk)
        g.skip_f32 = True
        nets.append((g, d))
    streams = [torch.cuda.Stream() for _ in range(nbuf)]

    def work():
        main = torch.cuda.current_stream()
        for s in streams:
            s.wait_stream(main)
        for c in range(P // chunk):
            k = c % nbuf
            with torch.cuda.stream(streams[k]):
                g, d = nets[k]
                g.forward(z[c * chunk:(c + 1) * chunk])
                d.forward(g.image_padded_bf16())
        for s in streams:
            main.wait_stream(s)

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        work()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        work()
    for _ in range(3):
        gr.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        gr.replay()
    e1.record()
    torch.cuda.synchronize()
    print(f"chunk={chunk:6d} streams={nbuf} G+D forward over {P} particles: {e0.elapsed_time(e1) / 10:.3f} ms", flush=True)
    del gr, nets
