import os, sys, subprocess, tempfile
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests'); sys.path.insert(0, '/root/repo/tools')
import sweep_9x45 as S
from test_workspace_scripts import h5pack
tmp = tempfile.mkdtemp()
x, lab = S.synth_set(1, 60000, 544)
path = os.path.join(tmp, "yosemite-m-unproj.h5")
h5pack(path, ("Label", lab.reshape(-1, 1), 128, 1), ("Distance", x, 128, 0))
os.makedirs(os.path.join(tmp, "out", "logging"))
r = subprocess.run([os.path.join('/root/repo', "bin", "pj-sweep"), path, os.path.join(tmp, "out"), "-iters", sys.argv[1], "-mu", S.MUS, "-gamma", S.GAMMAS, "-prox", "sign"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
print(r.stdout.decode()[-600:])
