import torch
n = 1 << 30
a = torch.empty(n, dtype=torch.bfloat16, device='cuda')
b = torch.empty(n, dtype=torch.bfloat16, device='cuda')
def t(f, it=10):
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it
ms = t(lambda: a.fill_(1.0)); print('fill (write only)  %.0f GB/s' % (n * 2 / ms / 1e6))
ms = t(lambda: a.zero_()); print('memset             %.0f GB/s' % (n * 2 / ms / 1e6))
ms = t(lambda: b.copy_(a)); print('copy (r+w)         %.0f GB/s' % (n * 4 / ms / 1e6))
ms = t(lambda: a.sum()); print('sum (read only)    %.0f GB/s' % (n * 2 / ms / 1e6))
# 1 read : 5 write mix like K4 (96 -> 510): y[5n] = f(x[n])
x = torch.empty(n // 4, dtype=torch.bfloat16, device='cuda')
y = torch.empty(5, n // 4, dtype=torch.bfloat16, device='cuda')
ms = t(lambda: torch.add(x.unsqueeze(0).expand(5, -1), 1.0, out=y)); print('1r:5w broadcast    %.0f GB/s' % (6 * (n // 4) * 2 / ms / 1e6))
