"""numpy model of the warp-level 2048-point real FFT used by the CUDA kernels (index algebra only).

One warp owns one frame.  The shifted frame x'[j] (live for j < 1102) is packed into
z[n] = x'[2n] + i x'[2n+1], n = n1 + 32*n2 with lane = n1 and register = n2, and transformed as
32 x 32:   pass 1 (over n2, in registers) -> twiddle W1024^(n1*k1) -> transpose (shared memory)
        -> pass 2 (over n1, in registers) -> Z[k1 + 32*k2] in lane k1 / register k2
        -> pairing with Z[1024-k] (lane (32-k1)%32, register 31-k2 (k1>0) or (32-k2)%32 (k1=0))
        -> real post-processing X[k] = (Z[k]+conj(Zp))/2 - i/2 * W2048^k * (Z[k]-conj(Zp)).
The inverse runs the same machinery on conj(Z).  This file checks that algebra against
numpy.fft so the CUDA transcription only has to follow it.  Run: python tools/fft_model.py
"""
import numpy as np

M = 1024
N = 2048


def fwd_model(xs):
    """xs: real[2048] (shifted frame). returns X[0..1024] via the lane/register algebra."""
    z = xs[0::2] + 1j * xs[1::2]
    v = np.zeros((32, 32), complex)                    # [lane n1][reg n2]
    for n1 in range(32):
        for n2 in range(32):
            v[n1, n2] = z[n1 + 32 * n2]
    W32 = np.exp(-2j * np.pi * np.outer(np.arange(32), np.arange(32)) / 32)
    A = v @ W32                                        # A[n1][k1] = sum_n2 v[n1][n2] W32^(n2 k1)
    tw = np.exp(-2j * np.pi * np.outer(np.arange(32), np.arange(32)) / 1024)
    A = A * tw                                         # W1024^(n1 k1)
    u = A.T                                            # transpose: lane k1 holds u[k1][n1]
    Zl = u @ W32                                       # Zl[k1][k2] = sum_n1 u[k1][n1] W32^(n1 k2)
    # pairing
    X = np.zeros(1025, complex)
    for k1 in range(32):
        for k2 in range(32):
            k = k1 + 32 * k2
            pl = (32 - k1) % 32
            pr = (31 - k2) if k1 else (32 - k2) % 32
            Zk, Zp = Zl[k1, k2], Zl[pl, pr]
            assert (pl + 32 * pr) % 1024 == (1024 - k) % 1024
            Wk = np.exp(-2j * np.pi * k / N)
            X[k] = 0.5 * (Zk + np.conj(Zp)) - 0.5j * Wk * (Zk - np.conj(Zp))
    Z0 = Zl[0, 0]
    X[1024] = Z0.real - Z0.imag
    return X


def inv_model(X):
    """X[0..1024] -> real[2048] via the same lane/register algebra (irfft semantics)."""
    Zl = np.zeros((32, 32), complex)
    for k1 in range(32):
        for k2 in range(32):
            k = k1 + 32 * k2
            kp = 1024 - k                              # partner bin (X[1024] for k = 0)
            Xe = 0.5 * (X[k] + np.conj(X[kp]))
            Xo = 0.5 * np.exp(+2j * np.pi * k / N) * (X[k] - np.conj(X[kp]))
            Zl[k1, k2] = Xe + 1j * Xo
    v = np.conj(Zl)                                    # IFFT(Z) = conj(FFT(conj Z)) / M
    W32 = np.exp(-2j * np.pi * np.outer(np.arange(32), np.arange(32)) / 32)
    tw = np.exp(-2j * np.pi * np.outer(np.arange(32), np.arange(32)) / 1024)
    A = (v @ W32) * tw
    zl = np.conj(A.T @ W32) / M                        # zl[lane n1'][reg n2'] = z[n1' + 32 n2']
    x = np.zeros(N)
    for n1 in range(32):
        for n2 in range(32):
            n = n1 + 32 * n2
            x[2 * n], x[2 * n + 1] = zl[n1, n2].real, zl[n1, n2].imag
    return x


def main():
    rng = np.random.RandomState(0)
    f = np.zeros(N)
    f[473:1575] = rng.randn(1102)                      # frame as the reference sees it (taps at 473..1574)
    xs = np.roll(f, -473)                              # shifted frame: live at 0..1101
    X_ref = np.fft.rfft(f)
    X = fwd_model(xs)
    k = np.arange(1025)
    print("forward |X| err", np.abs(np.abs(X) - np.abs(X_ref)).max())
    print("forward phase-corrected err", np.abs(X * np.exp(-2j * np.pi * 473 * k / N) - X_ref).max())
    x_back = inv_model(X)
    print("inverse err", np.abs(x_back - xs).max(), "tail", np.abs(x_back[1102:]).max())
    # GL step in the shifted domain equals the reference step: phase-normalise then inverse
    S = np.abs(rng.randn(1025)) + 0.1
    Xn_ref = S * X_ref / np.maximum(1e-8, np.abs(X_ref))
    f_ref = np.fft.irfft(Xn_ref)
    Xn = S * X / np.maximum(1e-8, np.abs(X))
    xs_new = inv_model(Xn)
    print("GL step shifted-domain err", np.abs(np.roll(xs_new, 473) - f_ref).max())


if __name__ == "__main__":
    main()
