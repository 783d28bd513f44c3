"""How many rows fall inside the 1xTF32 error window of the k-th best score? (design experiment)"""
import json, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import bench as B
dev = torch.device("cuda", 0)
C = 2.0**-10 + 2.0**-11 + 2.0**-14
for name in ("C2", "C1"):
    w = B.WORKLOADS[name]
    d = w["d"]
    xb_chunks = []
    xt, proj = B.gen_synthetic(torch, 1000, d, B.SEED_TRAIN, w["normalize"], dev)
    xq, _ = B.gen_synthetic(torch, 256, d, B.SEED_QUERY, w["normalize"], dev, proj=proj)
    gen = torch.Generator(device=dev); gen.manual_seed(B.SEED_BASE)
    nb, done = w["nb"], 0
    ip = w["metric"] == B.METRIC_IP
    allS = []
    ymax = 0.0
    while done < nb:
        ni = min(2_000_000, nb - done)
        z = torch.randn((ni, B.D_LATENT), generator=gen, device=dev, dtype=torch.float32)
        x = torch.sin(z @ proj)
        if w["normalize"]:
            x /= x.norm(dim=1, keepdim=True).clamp_min(1e-20)
        ymax = max(ymax, float(x.norm(dim=1).max()))
        if ip:
            s = -(xq.double() @ x.double().T)
        else:
            s = (xq.double()**2).sum(1, keepdim=True) - 2 * xq.double() @ x.double().T + (x.double()**2).sum(1)[None]
        allS.append(torch.topk(s, 2048, dim=1, largest=False).values)
        done += ni
    S = torch.sort(torch.cat(allS, 1), dim=1).values[:, :2048]
    xn = xq.norm(dim=1).double()
    E = (1.0 if ip else 2.0) * C * xn * ymax
    for kth in (10, 16):
        for mult in (1.0, 2.0):
            thr = S[:, kth - 1] + mult * E
            cnt = (S <= thr[:, None]).sum(1).float()
            print(json.dumps({"workload": name, "kth": kth, "window_mult_of_E": mult, "E_mean": float(E.mean()),
                              "score_k": float(S[:, kth - 1].mean()), "score_1": float(S[:, 0].mean()),
                              "count_mean": float(cnt.mean()), "count_p50": float(cnt.median()),
                              "count_p99": float(cnt.quantile(0.99)), "count_max": float(cnt.max())}), flush=True)
