import numpy as np, sys
a=np.fromfile(sys.argv[1],dtype=np.float32).reshape(64,96,3); b=np.fromfile(sys.argv[2],dtype=np.float32).reshape(64,96,3)
d=a-b
print(sys.argv[1].split('/')[-1], "relMSE %.3e"%float(np.mean(d**2/(b**2+1e-4))), "identical %.3f"%float((np.abs(d).max(2)==0).mean()), "within1pct %.3f"%float((np.abs(d).max(2)<=0.01*b.mean(2)+1e-3).mean()))
