import os, sys
sys.path.insert(0, '/root/repo')
from triton_client_b200 import _lib
_lib.LIB_PATH = os.path.join('/root/repo', 'profiles', 'probes', 'libtcb200_prof.so')   # built by profiles/probes/build_prof_lib.sh (post3d.cu with -DTCB_SWEEP_PROFILE)
exec(open('/root/repo/profiles/probes/prof_3d_rows.py').read().split("heads = synth")[0])
