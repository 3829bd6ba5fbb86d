import torch
n = 6 * (1 << 30) // 8
x = torch.empty(n, dtype=torch.float64, device="cuda"); y = torch.empty(n, dtype=torch.float64, device="cuda")
def t(f, reps=5):
    for _ in range(2): f()
    torch.cuda.synchronize()
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)/reps
ms = t(lambda: x.fill_(1.5)); print("fill 6 GiB: %.3f ms -> %.2f TB/s written" % (ms, 6*(1<<30)/ms/1e9))
ms = t(lambda: x.copy_(y)); print("copy 6 GiB: %.3f ms -> %.2f TB/s read+write" % (ms, 12*(1<<30)/ms/1e9))
ms = t(lambda: torch.cuda.memset if False else x.zero_()); print("zero 6 GiB: %.3f ms -> %.2f TB/s written" % (ms, 6*(1<<30)/ms/1e9))
