# Prototype of the register/smem Stockham scheme used by the CUDA kernels (index math check).
import numpy as np

def dft_small(x, inv):
    R = len(x)
    n = np.arange(R)
    s = 1 if inv else -1
    W = np.exp(s*2j*np.pi*np.outer(n, n)/R)
    return W @ x

def radices(N, EPT):
    rs = []
    rem = N
    while rem >= EPT:
        rs.append(EPT); rem //= EPT
    if rem > 1:
        rs.append(rem)
    return rs

def line_fft(x, EPT, inv=False):
    """x: length N complex. Emulates T=N/EPT threads each holding v[m]=x[t+m*T]."""
    N = len(x); T = N//EPT
    v = np.zeros((T, EPT), complex)
    for t in range(T):
        for m in range(EPT):
            v[t, m] = x[t+m*T]
    rs = radices(N, EPT)
    Ns = 1
    s = 1 if inv else -1
    for si, R in enumerate(rs):
        nb = EPT//R
        last = (si == len(rs)-1)
        sm = np.zeros(N, complex)
        for t in range(T):
            for b in range(nb):
                j = t + b*T
                k = j % Ns
                inp = np.array([v[t, b + r*nb] for r in range(R)])
                tw = np.exp(s*2j*np.pi*k*np.arange(R)/(Ns*R))
                out = dft_small(inp*tw, inv)
                j0 = (j//Ns)*Ns*R + k
                for r in range(R):
                    if last:
                        assert j0 + r*Ns == t + (b + r*nb)*T
                        v[t, b + r*nb] = out[r]
                    else:
                        sm[j0 + r*Ns] = out[r]
        if not last:
            for t in range(T):
                for m in range(EPT):
                    v[t, m] = sm[t+m*T]
        Ns *= R
    y = np.zeros(N, complex)
    for t in range(T):
        for m in range(EPT):
            y[t+m*T] = v[t, m]
    return y

rng = np.random.default_rng(0)
for N, EPT in [(16,8),(32,8),(64,8),(128,8),(256,8),(32,16),(64,16),(128,16),(256,16),(512,16),(16,16),(8,8),(512,8)]:
    x = rng.normal(size=N)+1j*rng.normal(size=N)
    e1 = np.abs(line_fft(x,EPT)-np.fft.fft(x)).max()
    e2 = np.abs(line_fft(x,EPT,True)-np.fft.ifft(x)*N).max()
    print(N,EPT,radices(N,EPT),e1,e2)

# R2C untangle / C2R retangle with packed column 0, scale check
BX=128; H=BX//2
x = rng.normal(size=BX)
z = x[0::2]+1j*x[1::2]
Z = np.fft.fft(z)
k = np.arange(H)
Zp = np.conj(Z[(H-k)%H])
w = np.exp(-2j*np.pi*k/BX)
X2 = (Z+Zp) - 1j*w*(Z-Zp)       # = 2*X[k], k<H
XH2 = 2*(Z[0].real - Z[0].imag) # 2*X[H]
Xref = np.fft.rfft(x)
print('r2c', np.abs(X2-2*Xref[:H]).max(), abs(XH2-2*Xref[H]))
# inverse: from X (k=0..H) -> Z' = (X[k]+conj(X[H-k])) + i*conj(w^k)*(X[k]-conj(X[H-k]))
Xf = Xref
Xk = Xf[:H]; Xp = np.conj(Xf[H-k])
Zi = (Xk+Xp) + 1j*np.conj(w)*(Xk-Xp)
zi = np.fft.ifft(Zi)*H
xr = np.empty(BX); xr[0::2]=zi.real; xr[1::2]=zi.imag
print('c2r scale', np.abs(xr - BX*x).max())
