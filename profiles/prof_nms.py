import ctypes, torch, numpy as np, sys
sys.path.insert(0,'.')
from triton_client_b200 import synth, _lib
from triton_client_b200.gpu import Engine
import os
_lib.LIB_PATH=os.environ.get('TCB_PROF_LIB') or os.path.join('profiles','probes','libtcb200_prof.so')   # built by profiles/probes/build_prof_lib.sh (yolo_post.cu with -DTCB_NMS_PROFILE)
eng=Engine(0)
pred=torch.from_numpy(synth.yolo_output(64,25200,2,2002)).cuda()
l=_lib.lib()
buf=(ctypes.c_longlong*16)()
cta=(ctypes.c_longlong*4096)()
for it in range(3):
    eng.yolo_nms(pred, conf_thres=0.3); torch.cuda.synchronize()
    l.tcb_debug_nms_prof(buf, 1)
    v=list(buf)
    print('sub-phases: chunk table+scan',v[8],'gather',v[9],'fill (in load)',v[0],'| rows written',v[10],'fence+ticket',v[11],'merge (in write)',v[6])
    print('cycles: load',v[0],'sort',v[1],'roundload',v[2],'phase1+compact',v[3],'phase2',v[4],'sweep',v[5],'write',v[6],'kept*1e6+n',v[7])
l.tcb_debug_nms_cta(cta)
a=np.array(list(cta)).reshape(1024,4)[:128]
print('per CTA (128): cycles mean %.0f max %d min %d | n mean %.0f max %d | rounds mean %.2f max %d | kept mean %.1f max %d' % (a[:,0].mean(), a[:,0].max(), a[:,0].min(), a[:,1].mean(), a[:,1].max(), a[:,2].mean(), a[:,2].max(), a[:,3].mean(), a[:,3].max()))
o=np.argsort(-a[:,0])[:8]
print('slowest:', [(int(i), a[i].tolist()) for i in o])
print('corr cycles~rounds', np.corrcoef(a[:,0],a[:,2])[0,1], 'cycles~n', np.corrcoef(a[:,0],a[:,1])[0,1])
