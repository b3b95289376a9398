import sys, torch, statistics
sys.path.insert(0, '/root/repo')
from distributed_training_b200.ops import native
B,H,T=8,12,1024; C=H*64
qkv=torch.randn(B,T,3*C,device='cuda').to(torch.bfloat16); dout=torch.randn(B,T,C,device='cuda').to(torch.bfloat16)
out=torch.empty(B,T,C,device='cuda',dtype=torch.bfloat16); lse=torch.empty(B,H,T,device='cuda'); dqkv=torch.empty_like(qkv)
scr=torch.empty(native.attention_bwd_scratch_floats(B,T,H),device='cuda')
native.attention_fwd(qkv,out,lse,H)
fl=torch.empty(256<<20,dtype=torch.uint8,device='cuda')
ts=[]
for i in range(15):
    fl.zero_(); s,e=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    s.record(); native.attention_bwd(qkv,out,dout,lse,dqkv,scr,H); e.record(); torch.cuda.synchronize(); ts.append(s.elapsed_time(e)*1e3)
print(sys.argv[1] if len(sys.argv)>1 else '', round(statistics.median(ts[5:]),1))
