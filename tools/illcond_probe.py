import sys; sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import numpy as np, torch
from metaopt_b200.engine import get_engine
from oracle import gp as ogp
rs = np.random.RandomState(3)
centres = rs.uniform(size=(20, 3))
X = np.concatenate([c + 1e-3 * rs.randn(10, 3) for c in centres])
y = np.sin(3 * X).sum(axis=1)
C = rs.uniform(size=(500, 3))
eng = get_engine(0)
eng.fit(X, y, alpha=1e-10, normalize_y=True, length_scale=0.5)
ref = ogp.SklearnGP(0, 1.0, 0.5 * np.ones(3), 1e-10, True).fit(X, y)
mu_r, sd_r = ref.predict(C)
out = eng.score(torch.as_tensor(C).cuda(), acq_mask=1, y_opt=float(y.min()), dense=True)
mu, sd = out["mu"].cpu().numpy(), out["sigma"].cpu().numpy()
# long double reference
ld = np.longdouble
Xs = (X/0.5).astype(ld); Cs=(C/0.5).astype(ld)
def kern(A,B):
    d2=((A[:,None,:]-B[None,:,:])**2).sum(-1); k=np.sqrt(d2)*np.sqrt(ld(5)); return (1+k+k*k/3)*np.exp(-k)
K=kern(Xs,Xs); K[np.diag_indices_from(K)]+=ld(1e-10)
n=len(X); L=np.zeros((n,n),dtype=ld)
for j in range(n):
    s=K[j,j]-(L[j,:j]**2).sum(); L[j,j]=np.sqrt(s)
    for i in range(j+1,n): L[i,j]=(K[i,j]-(L[i,:j]*L[j,:j]).sum())/L[j,j]
ym=y.mean(); ys=y.std(); yn=((y-ym)/ys).astype(ld)
def fsub(L,b):
    x=np.zeros_like(b)
    for i in range(len(b)): x[i]=(b[i]-(L[i,:i]*x[:i]).sum())/L[i,i]
    return x
Ks=kern(Cs,Xs)
V=np.array([fsub(L,Ks[i]) for i in range(len(C))])
var=(1-(V**2).sum(1)); var[var<0]=0; sd_x=(np.sqrt(var)*ys).astype(float)
print('cond', np.linalg.cond(K.astype(float)))
print('max |sd_gpu - sd_exact|', np.abs(sd-sd_x).max(), ' max |sd_sklearn - sd_exact|', np.abs(sd_r-sd_x).max(), ' max|gpu-sk|', np.abs(sd-sd_r).max())
print('var abs: gpu', np.abs(sd**2-sd_x**2).max(), 'sk', np.abs(sd_r**2-sd_x**2).max(), 'ystd2', ys**2)
