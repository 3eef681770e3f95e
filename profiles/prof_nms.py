import ctypes, torch, numpy as np, sys
sys.path.insert(0,'.')
from triton_client_b200 import synth, _lib
from triton_client_b200.gpu import Engine
import os
_lib.LIB_PATH=os.path.join('profiles','probes','libtcb200_prof.so')   # built by profiles/probes/build_prof_lib.sh (yolo_post.cu with -DTCB_NMS_PROFILE)
eng=Engine(0)
pred=torch.from_numpy(synth.yolo_output(64,25200,2,2002)).cuda()
l=_lib.lib()
buf=(ctypes.c_longlong*8)()
for it in range(3):
    eng.yolo_nms(pred, conf_thres=0.3); torch.cuda.synchronize()
    l.tcb_debug_nms_prof(buf, 1)
    v=list(buf)
    print('cycles: load',v[0],'sort',v[1],'roundload',v[2],'phase1+compact',v[3],'phase2',v[4],'sweep',v[5],'write',v[6],'kept*1e6+n',v[7])
