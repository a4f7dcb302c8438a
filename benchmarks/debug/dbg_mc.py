import sys, torch, numpy as np
sys.path.insert(0, '.')
from erasurehead_b200.ops import mc, synth
torch.manual_seed(0)
rows, d, C, model = 256, 128, 4, 2
ld = synth.pad_cols(d)
X = torch.zeros(rows, ld, device='cuda'); X[:, :d] = torch.randn(rows, d, device='cuda') / np.sqrt(d)
y = torch.randint(0, C, (rows,), device='cuda').float()
W = torch.randn(d, C, device='cuda'); Wp = mc.pack_params(W)
G = torch.full((16, ld), 7.0, device='cuda')
loss = torch.zeros((), dtype=torch.float64, device='cuda')
scr = mc.scratch(rows, d, C, 'cuda')
job = mc.make_job([X], [y], Wp.data_ptr(), G.data_ptr(), d, C, [1.0], model, loss_ptr=loss.data_ptr())
mc.launch(job, scr); torch.cuda.synchronize()
P = X[:, :d].double() @ W.double(); S = torch.softmax(P, 1)
print('loss', loss.item(), 'want', -torch.log(S[torch.arange(rows), y.long()]).sum().item())
S[torch.arange(rows), y.long()] -= 1
f = scr.view(torch.float32)
ldr = (rows + 31)//32*32
Rt = f[:16*ldr].reshape(16, ldr)
print('Rt err', (Rt[:C, :rows].double() - S.t()).abs().max().item(), 'Rt pad max', Rt[C:].abs().max().item())
Gs = f[16*ldr:16*ldr+16*ld].reshape(16, ld)
want = (X[:, :d].double().t() @ S).t()
print('G scratch max', Gs.abs().max().item(), 'G out max', G.abs().max().item(), 'want max', want.abs().max().item())
print('G err', (Gs[:C, :d].double() - want).abs().max().item())
print(Gs[:2, :8], want[:2, :8])
