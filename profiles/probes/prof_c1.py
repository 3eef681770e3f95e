"""Workload for an ncu launch list of configs[0] (YOLOv5n COCO, batch 1): image_adjust + decode/NMS of 1x25200x85."""
import sys
import torch
sys.path.insert(0, '/root/repo')
from triton_client_b200 import synth
from triton_client_b200.gpu import Engine
eng = Engine(0, workspace_bytes=512 << 20)
img = torch.from_numpy(synth.image_noise(640, 640, 1)[None]).cuda()
pred = torch.from_numpy(synth.yolo_output(1, 25200, 80, 5)).cuda()
for _ in range(4):
    eng.preprocess_images(img, 640, 640)
    eng.yolo_nms(pred, conf_thres=0.3)
torch.cuda.synchronize()
