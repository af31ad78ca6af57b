# PNG decode rate of cvr_decode_png_batch against the number of files per call:  python profiles/side_png_scale.py 220 2200 [--pageable]
import importlib, sys, json, time, numpy as np, torch, cv2
sys.path.insert(0,'.')
import bench
cvr = importlib.import_module("master-thesis_b200")
torch.cuda.set_device(0); dev=torch.device("cuda",0)
Ls, Rs = bench.make_pairs(8)
files = [cv2.imencode(".png", im)[1].tobytes() for im in list(Ls) + list(Rs)]
for n in [int(a) for a in sys.argv[1:] if not a.startswith('--')]:
    sizes = [len(files[i % len(files)]) for i in range(n)]
    pin = torch.empty(int(np.sum(sizes)), dtype=torch.uint8)
    if "--pageable" not in sys.argv: pin = pin.pin_memory()
    arr = pin.numpy(); batch = []; o = 0
    for i in range(n):
        arr[o:o + sizes[i]] = np.frombuffer(files[i % len(files)], np.uint8); batch.append(arr[o:o + sizes[i]]); o += sizes[i]
    rec = cvr.Reconstruct3D(bench.W, bench.H, filter_only=True, max_points=16, device=0)
    out = torch.empty((n, bench.H, bench.W, 3), dtype=torch.uint8, device=dev)
    rec.DecodePNG(batch, device_out=out.data_ptr())
    torch.cuda.synchronize(); t0=time.perf_counter()
    for _ in range(3): rec.DecodePNG(batch, device_out=out.data_ptr())
    torch.cuda.synchronize(); dt=time.perf_counter()-t0
    rec.set_profiling(True); t1=time.perf_counter(); rec.DecodePNG(batch, device_out=out.data_ptr()); t2=time.perf_counter()
    print(n, round(3*n/dt,1), 'images/s;  one call %.1f ms, stages (ms):' % ((t2-t1)*1e3), {k: round(v,1) for k,v in rec.last_timings()}, flush=True)
    rec.close()
