import numpy as np, sys
p=np.load('/tmp/ad_precise.npy'); a=np.load('/tmp/ad_tc5.npy'); b=np.load('/tmp/ad_mma.npy')
print('tc5 vs precise max|d|', float(np.abs(a-p).max()), ' mma vs precise', float(np.abs(b-p).max()), ' tc5 vs mma', float(np.abs(a-b).max()), 'nan', int(np.isnan(a).sum()))
