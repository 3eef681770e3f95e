import torch, sys
sys.path.insert(0,'.')
from triton_client_b200.gpu import Engine
def t(fn, it=30):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a,b=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    ts=[]
    for _ in range(it):
        a.record(); fn(); b.record(); b.synchronize(); ts.append(a.elapsed_time(b))
    ts.sort(); return ts[len(ts)//2]
o=torch.empty((64,3,640,640),device='cuda'); o2=torch.empty_like(o)
ms=t(lambda:o.fill_(1.0)); print('fill 314MB', ms, 'ms', 314.57/ms, 'GB/s')
ms=t(lambda:o2.copy_(o)); print('copy 314MB->314MB', ms, 'ms', 2*314.57/ms, 'GB/s')
src=torch.randint(0,255,(64,720,1280,3),dtype=torch.uint8,device='cuda')
s2=torch.empty_like(src)
ms=t(lambda:s2.copy_(src)); print('copy u8 177MB', ms, 2*176.9/ms)
eng=Engine(0)
tiny=torch.zeros((64,8,8,3),dtype=torch.uint8,device='cuda')
ms=t(lambda:eng.preprocess_images(tiny,640,640,mode='pad',out=o)); print('preprocess pad-only (pure write 314MB)', ms, 314.57/ms)
ms=t(lambda:eng.preprocess_images(src,640,640,mode='letterbox',out=o)); print('letterbox', ms, 491.52/ms)
ms=t(lambda:eng.preprocess_images(src,640,640,mode='stretch',out=o)); print('stretch', ms, 491.52/ms)
big=torch.empty((4,64,3,640,640),device='cuda')
ms=t(lambda:big.fill_(1.0)); print('fill 1.26GB', ms, 4*314.57/ms)
