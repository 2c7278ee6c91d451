import importlib, sys, numpy as np
sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import helpers as H
svdj = importlib.import_module("svd-jacobi-mpi-cuda_b200")
m=n=512
A0=H.uniform_matrix(m,n,seed=1)
A=A0.copy(order="F"); s=np.zeros(n); V=np.zeros((n,n),order="F")
try:
    st=svdj.dgesvd_mg(0,0,m,n,A,m,s,V,n,2,[0,0])
    print("ok sweeps",st.sweeps, H.sigma_error_normwise(s,np.linalg.svd(A0,compute_uv=False)), H.svd_quality(A0,A,s,V))
except Exception as e:
    print("FAIL", e)
