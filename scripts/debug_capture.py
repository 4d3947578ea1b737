import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import distributed_tensorflow_framework_b200 as dtf
from distributed_tensorflow_framework_b200 import models
from distributed_tensorflow_framework_b200.ops import _lib
from distributed_tensorflow_framework_b200.engine import variables as V
from distributed_tensorflow_framework_b200.engine.tower import Tower

import threading
orig = _lib.call
def traced(name, *args):
    sp = _lib.stream_ptr()
    st0 = _lib.lib().dtf_stream_capture_status(sp)
    try:
        rc = orig(name, *args)
    except Exception as e:
        print("  call %s raised %s" % (name, str(e)[:80]), flush=True); raise
    st1 = _lib.lib().dtf_stream_capture_status(sp)
    if TR["on"]:
        print("  %-26s thread %s stream %x status %d->%d" % (name, threading.current_thread().name, sp, st0, st1), flush=True)
    return rc
TR = {"on": False}
_lib.call = traced

def status(tag):
    st = _lib.lib().dtf_stream_capture_status(_lib.stream_ptr())
    print("%-30s status %d" % (tag, st), flush=True)

x = torch.rand(64, 32, 32, 3).cuda(); y = torch.randint(0, 10, (64,)).cuda()
tr = dtf.Trainer(dtf.Net(models.Cifar10CNN()), dtf.SoftmaxCrossEntropyLoss(), dtf.MomentumOptimizer(0.02, 0.9), device="cuda:0")
side = torch.cuda.Stream()
side.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(side):
    for _ in range(3):
        tr.step_device([(x, y)])
torch.cuda.current_stream().wait_stream(side)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
mode = sys.argv[1] if len(sys.argv) > 1 else "thread_local"
same = len(sys.argv) > 2
import traceback
try:
    with torch.cuda.graph(g, capture_error_mode=mode, **({"stream": side} if same else {})):
        status("begin")
        tr.engine.begin_step(10)
        status("after begin_step")
        with V.using_store(tr.store):
            tower = Tower(tr.net, "tower_0", [], x, y, tr.loss, tr.optimizer, tr.store)
            loss, logits = tower.tower_loss(None)
            status("after forward")
            TR["on"] = False
            loss.backward()
            TR["on"] = False
            status("after backward")
        tr.engine.finish_step()
        status("after finish")
    print("capture ok")
    g.replay(); torch.cuda.synchronize(); print("replay ok", float(loss))
except Exception as e:
    print("ERR", str(e)[:120]); traceback.print_exc(limit=6)
