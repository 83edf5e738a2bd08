import torch, sys
sys.path.insert(0, '.')
from llm_in_practise_b200.ops import gemm as G
torch.manual_seed(0)
for (M,N,K) in [(128,64,32),(256,128,64),(512,768,264)]:
    x = torch.randn(M,K,device='cuda'); w = torch.randn(N,K,device='cuda'); dy = torch.randn(M,N,device='cuda')
    for name, fn, ref in [('fwd', lambda o: G.gemm(x,w,out=o), x@w.t()), ('dgrad', lambda o: G.linear_dgrad(dy,w,out=o), dy@w), ('wgrad', lambda o: G.linear_wgrad(dy,x,out=o), dy.t()@x)]:
        out = torch.full(ref.shape, float('nan'), device='cuda')
        fn(out); torch.cuda.synchronize()
        print(M,N,K,name,'nan',out.isnan().float().mean().item(),'zero',(out==0).float().mean().item(),'relerr',((out-ref).abs().max()/ref.abs().max()).item())
        if name!='fwd': print(out[:2,:5].tolist(), ref[:2,:5].tolist())
