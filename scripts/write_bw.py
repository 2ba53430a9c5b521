"""Write-only and copy bandwidth of this B200 (what a store-bound kernel can hope for)."""
import torch, json
dev = torch.device("cuda:0")
n = 1 << 31                      # 2 GiB
a = torch.empty(n, dtype=torch.uint8, device=dev)
b = torch.empty(n, dtype=torch.uint8, device=dev)
def t(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
ms_fill = t(lambda: a.zero_())
ms_copy = t(lambda: b.copy_(a))
a16 = a.view(torch.int16); 
ms_fill1 = t(lambda: a.fill_(1))
print(json.dumps({"memset_GBs": n / ms_fill / 1e6, "fill1_GBs": n / ms_fill1 / 1e6, "copy_GBs_rw": 2 * n / ms_copy / 1e6}))
