import sys, os
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import config_sweep as cs
for n, ns in ((256, 2), (1024, 2), (512, 4)):
    cs.c5(n, ns, 1)
