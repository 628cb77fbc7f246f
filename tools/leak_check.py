import importlib, sys, os, numpy as np, torch
sys.path.insert(0, os.getcwd())
orbx = importlib.import_module("3_orb_slam3_selfnote_b200"); synth = importlib.import_module("3_orb_slam3_selfnote_b200.synth")
img = synth.frame(752, 480, 7, 0); img2 = synth.frame(640, 480, 7, 1)
torch.cuda.init()
def used(): 
    free, total = torch.cuda.mem_get_info(); return (total - free) / 1e6
base = None
for it in range(150):
    gx = orbx.ORBextractor(1000)
    gx(img, None, (0, 0)); gx(img2, None, (0, 0)); gx(img, None, (0, 0))
    m = orbx.ORBmatcher()
    gx.close(); m.close()
    if it == 10: base = used()
    if it % 50 == 49: print(it, round(used(), 1), "MB (base", round(base, 1), ")")
import resource
print("host maxrss MB", resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 1024)
