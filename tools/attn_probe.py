import torch, torch.nn.functional as F
from torch.profiler import profile, ProfilerActivity
B,T,H,dh=32,1024,12,64
d=H*dh
qkv=torch.randn(B*T,3*d,device='cuda',dtype=torch.bfloat16)
v5=qkv.view(B,T,3,H,dh)
q,k,v=(v5[:,:,i].permute(0,2,1,3).detach().requires_grad_(True) for i in range(3))
print('q strides',q.stride(), q.is_contiguous())
with torch.enable_grad():
    o=F.scaled_dot_product_attention(q,k,v,is_causal=True)
print('o shape',o.shape,'strides',o.stride())
a=o.detach().permute(0,2,1,3); print('a contiguous',a.is_contiguous())
da=torch.randn(B*T,d,device='cuda',dtype=torch.bfloat16)
go=da.view(B,T,H,dh).permute(0,2,1,3)
dq,dk,dv=torch.autograd.grad(o,(q,k,v),go,retain_graph=True)
print('dq strides',dq.stride(),'dk',dk.stride(),'dv',dv.stride())
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    with torch.enable_grad():
        o=F.scaled_dot_product_attention(q,k,v,is_causal=True)
    torch.cuda.synchronize()
    dq,dk,dv=torch.autograd.grad(o,(q,k,v),go)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=12, max_name_column_width=70))
# separate q,k,v contiguous [B,T,H,dh]
qs,ks,vs=(torch.randn(B,T,H,dh,device='cuda',dtype=torch.bfloat16).transpose(1,2).requires_grad_(True) for _ in range(3))
with torch.enable_grad():
    o2=F.scaled_dot_product_attention(qs,ks,vs,is_causal=True)
print('o2 strides',o2.stride())
g=torch.autograd.grad(o2,(qs,ks,vs),go)
print('sep dq strides',g[0].stride())
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    with torch.enable_grad():
        o2=F.scaled_dot_product_attention(qs,ks,vs,is_causal=True)
    g=torch.autograd.grad(o2,(qs,ks,vs),go)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=12, max_name_column_width=70))
