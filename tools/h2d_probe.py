import torch, time
dev=torch.device("cuda",0)
def run(chunk_mb, n_chunks, tiny_pageable=False, reps=20):
    hs=[torch.empty(int(chunk_mb*1e6), dtype=torch.uint8).pin_memory() for _ in range(n_chunks)]
    ds=[torch.empty(int(chunk_mb*1e6), dtype=torch.uint8, device=dev) for _ in range(n_chunks)]
    tiny_h=[torch.empty(4096, dtype=torch.uint8) for _ in range(n_chunks)]
    tiny_d=[torch.empty(4096, dtype=torch.uint8, device=dev) for _ in range(n_chunks)]
    s=torch.cuda.Stream()
    torch.cuda.synchronize()
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s):
        for _ in range(3):
            for h,d in zip(hs,ds): d.copy_(h, non_blocking=True)
        e0.record()
        for _ in range(reps):
            for i,(h,d) in enumerate(zip(hs,ds)):
                d.copy_(h, non_blocking=True)
                if tiny_pageable: tiny_d[i].copy_(tiny_h[i], non_blocking=True)
        e1.record()
    torch.cuda.synchronize()
    ms=e0.elapsed_time(e1)
    print(f"{n_chunks} x {chunk_mb} MB{' + tiny pageable' if tiny_pageable else ''}: {chunk_mb*n_chunks*reps/ms:.1f} GB/s")
run(66.7,1); run(9.5,7); run(9.5,7,True); run(38,7); run(2.4,7); run(2.4,7,True)
