import itertools, math
def wavefronts(N, threads, P1, P2, SH):
    NN=N*N; EPB=threads//NN
    tot={'T':0,'S':0,'R':0}; ideal=0
    for w0 in range(0, threads, 16):
        lanes=[t for t in range(w0, min(w0+16, threads)) if t//NN < EPB]
        if not lanes: continue
        for o in 'TSR':
            for i in range(N):
                banks={}
                for t in lanes:
                    slot, lt = divmod(t, NN); lx, ly = lt % N, lt // N
                    if o=='T': idx = lx + P1*ly + P2*i
                    elif o=='S': idx = lx + P1*i + P2*ly
                    else: idx = i + P1*lx + P2*ly
                    idx += slot*SH
                    banks.setdefault(idx % 16, set()).add(idx)
                tot[o] += max(len(v) for v in banks.values())
        ideal += N
    return tot, ideal
for N, threads in [(3,64),(5,128),(6,128),(7,64),(2,32)]:
    NN=N*N; N3=N**3
    base_SH = 3*N3 + 12*NN + 12
    t0,ideal = wavefronts(N, threads, N, NN, base_SH)
    best=None
    for P1 in range(N, N+5):
        for P2 in range(P1*N, P1*N+17):
            Vsz = P2*(N-1) + P1*(N-1) + N   # max index + 1
            for shpad in range(0, 16):
                SH = 3*Vsz + 12*NN + 12 + shpad
                t,_ = wavefronts(N, threads, P1, P2, SH)
                s = sum(t.values())
                if best is None or (s, SH) < (best[0], best[4]): best=(s, P1, P2, t, SH, Vsz)
    print(N, threads, 'base', t0, sum(t0.values()), 'ideal', 3*ideal, 'best', best)
