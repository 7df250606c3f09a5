"""Cost of the exact nearest-neighbour stand-in for the un-vendored knn_cuda (oracle/ref_py.py BruteForceKNN) that the
reference arm of bench.py calls once per view, next to its ingredients.  python tools/knn_standin_time.py"""
import sys, torch, time
sys.path.insert(0,'/root/repo')
from oracle.ref_py import BruteForceKNN
dev='cuda'
torch.manual_seed(0)
ref=torch.rand(1,6890,3,device=dev); q=ref[:, torch.randint(0,6890,(50000,),device=dev)]+0.01*torch.randn(1,50000,3,device=dev)
def t(fn,n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); a=torch.cuda.Event(enable_timing=True); b=torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize(); return a.elapsed_time(b)/n
for k in (1, 2, 3):
    print('standin k',k, round(t(lambda: BruteForceKNN(k)(ref,q)),3),'ms')
print('cdist mm only', round(t(lambda: torch.cdist(q,ref)),3))
d=torch.cdist(q,ref)
print('topk8', round(t(lambda: torch.topk(d,8,dim=-1,largest=False)),3))
print('min', round(t(lambda: d.min(-1)),3))
print('exact cdist', round(t(lambda: torch.cdist(q,ref,compute_mode="donot_use_mm_for_euclid_dist"),n=3),3))
