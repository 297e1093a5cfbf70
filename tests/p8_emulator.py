"""Test-only float64 interpreter of the plan's layer program (mtldas_plan_describe JSON).

It executes exactly the contract the CUDA kernels implement -- P8 / P8S2 layouts, shifted-tap
GEMMs over flattened padded positions, packed weight placement, fused epilogues, pad zeroing --
with torch CPU ops, so the host-side planner (offsets, guards, phases, chunk slices, tap shifts)
is verified against the oracle on a machine without a GPU.  Not part of the product.
"""
import math

import torch

from oracle import mtl_oracle as O


class Emu:
    def __init__(self, desc, sd, nb):
        self.d, self.nb = desc, nb
        self.sd = {k: v.double() for k, v in sd.items() if v.is_floating_point()}
        self.layers = {L["conv"]: L for L in desc["layers"]}
        self.t = {}
        for t in desc["tensors"]:
            self.t[t["name"]] = dict(info=t, data=torch.zeros(t["nphase"], t["nchunks"], t["plane"], 8, dtype=torch.float64))
            assert t["g_lo"] >= t["Wp"] + 1
            assert t["g_lo"] + desc["micro_batch"] * t["Pimg"] + 512 + t["Wp"] + 1 <= t["plane"]

    # ---- folded weights of one reference layer ------------------------------------------------
    def folded(self, conv):
        L = self.layers[conv]
        spec = dict(conv=L["conv"], bn=L["bn"], bias=bool(L["bias"]), cout=L["cout"])
        return O.fold_bn_eval(self.sd, spec)

    def packed(self, op, g):
        ntaps, K, Cout = len(op["taps"]), op["nchunks"] * 8, op["Cout"]
        Wp = torch.zeros(ntaps, K, Cout, dtype=torch.float64)
        bias = torch.zeros(Cout, dtype=torch.float64)
        for s in g["segs"]:
            w, b = self.folded(s["layer"])
            if s.get("co_n", 0) > 0:        # N-split group: a slice of the layer's output channels
                w, b = w[s["co_lo"]:s["co_lo"] + s["co_n"]], b[s["co_lo"]:s["co_lo"] + s["co_n"]]
            co, ci, k, _ = w.shape
            for c in range(ci):
                shift = 1 if (s["split"] >= 0 and c >= s["split"]) else 0
                kk = (c - s["split"] if shift else c) + s["k_off"]
                for r in range(k):
                    for q in range(k):
                        Wp[r * k + q + shift, kk, s["cout_off"]:s["cout_off"] + co] = w[:, c, r, q]
            bias[s["cout_off"]:s["cout_off"] + co] = b
        return Wp, bias

    # ---- view addressing ----------------------------------------------------------------------
    def coords(self, H, W):
        Wp, Pimg = W + 1, (H + 1) * (W + 1)
        f = torch.arange(self.nb * Pimg)
        n, rem = f // Pimg, f % Pimg
        h, w = rem // Wp, rem % Wp
        return f, n, h, w, (h < H) & (w < W)

    def gather(self, name, chunk, f, n, h, w):
        t = self.t[name]
        i = t["info"]
        if i["layout"] == 0:
            return t["data"][0, chunk, i["g_lo"] + f]
        ph = (h & 1) * 2 + (w & 1)
        pos = i["g_lo"] + n * i["Pimg"] + (h >> 1) * i["Wp"] + (w >> 1)
        return t["data"][ph, chunk, pos]

    def scatter(self, name, chunk, f, n, h, w, valid, vals):
        t = self.t[name]
        i = t["info"]
        if i["layout"] == 0:
            t["data"][0, chunk, i["g_lo"] + f] = torch.where(valid[:, None], vals, torch.zeros_like(vals))
        else:
            ph = (h & 1) * 2 + (w & 1)
            pos = i["g_lo"] + n * i["Pimg"] + (h >> 1) * i["Wp"] + (w >> 1)
            t["data"][ph[valid], chunk, pos[valid]] = vals[valid]

    # ---- ops ----------------------------------------------------------------------------------
    def stem(self, op, x):
        w, b = self.folded(op["layer"])
        y = torch.relu(torch.nn.functional.conv2d(x.double(), w, b, stride=3, padding=2))
        self.write_nchw(op["out"], y)

    def write_nchw(self, name, y):
        i = self.t[name]["info"]
        f, n, h, w, valid = self.coords(i["H"], i["W"])
        for c in range(i["nchunks"]):
            vals = torch.zeros(len(f), 8, dtype=torch.float64)
            vals[valid] = y[n[valid], c * 8:(c + 1) * 8, h[valid], w[valid]]
            self.scatter(name, c, f, n, h, w, valid, vals)

    def read_nchw(self, name, chunk0=0, nch=None):
        i = self.t[name]["info"]
        nch = i["nchunks"] - chunk0 if nch is None else nch
        f, n, h, w, valid = self.coords(i["H"], i["W"])
        y = torch.zeros(self.nb, nch * 8, i["H"], i["W"], dtype=torch.float64)
        for c in range(nch):
            v = self.gather(name, chunk0 + c, f, n, h, w)
            y[n[valid], c * 8:(c + 1) * 8, h[valid], w[valid]] = v[valid]
        return y

    def conv(self, op):
        H, W = op["H"], op["W"]
        f, n, h, w, valid = self.coords(H, W)
        body = len(f)
        tile = op["MT"] * 128
        for g in op["groups"]:
            Wp, bias = self.packed(op, g)
            acc = torch.zeros(body, op["Cout"], dtype=torch.float64)
            for ti, (s, off) in enumerate(op["taps"]):
                name, chunk0, phase = g["src"][s]
                t = self.t[name]
                i = t["info"]
                assert i["Wp"] == W + 1 and i["Pimg"] == (H + 1) * (W + 1), (op["name"], name)
                # slab range safety for every CTA tile of the real kernel
                ntiles = (body + tile - 1) // tile
                assert i["g_lo"] - op["halo_lo"] >= 0
                assert i["g_lo"] + ntiles * tile + op["halo_hi"] <= i["plane"]
                assert -op["halo_lo"] <= off <= op["halo_hi"]
                ph = max(phase, 0)
                X = t["data"][ph, chunk0:chunk0 + op["nchunks"], i["g_lo"] + off:i["g_lo"] + off + body]  # [c, body, 8]
                acc += X.permute(1, 0, 2).reshape(body, -1) @ Wp[ti]
            y = acc + bias
            if op["epi"] in (2, 3):
                an, ac0, anch = g["aux"]
                a = torch.cat([self.gather(an, ac0 + (c % anch), f, n, h, w) for c in range(op["Cout"] // 8)], dim=1)
                y = torch.relu(y + a) if op["epi"] == 2 else torch.sigmoid(y) * a
            elif op["epi"] == 1:
                y = torch.relu(y)
            on, oc0 = g["out"]
            if "pool" in g:      # fused ReLU + MaxPool2d(2,2,ceil): only the pooled tensor is written
                full = torch.zeros(self.nb, op["Cout"], H, W, dtype=torch.float64)
                full[n[valid], :, h[valid], w[valid]] = y[valid]
                pooled = torch.nn.functional.max_pool2d(full, 2, 2, ceil_mode=True)
                pi = self.t[g["pool"]]["info"]
                fo, no, ho, wo, vo = self.coords(pi["H"], pi["W"])
                for c in range(op["Cout"] // 8):
                    vals = torch.zeros(len(fo), 8, dtype=torch.float64)
                    vals[vo] = pooled[no[vo], c * 8:(c + 1) * 8, ho[vo], wo[vo]]
                    self.scatter(g["pool"], oc0 + c, fo, no, ho, wo, vo, vals)
                continue
            for c in range(op["Cout"] // 8):
                self.scatter(on, oc0 + c, f, n, h, w, valid, y[:, c * 8:(c + 1) * 8])

    def pool(self, op):
        y = self.read_nchw(op["in"])
        self.write_nchw(op["out"], torch.nn.functional.max_pool2d(y, 2, 2, ceil_mode=True))

    def head(self, op):
        i = self.t[op["in"]]["info"]
        blk = self.t[op["in"]]["data"][0, op["chunk0"]:op["chunk0"] + 16, i["g_lo"]:i["g_lo"] + self.nb * i["Pimg"]]
        s = blk.reshape(16, self.nb, i["Pimg"], 8).sum(2).permute(1, 0, 2).reshape(self.nb, 128)
        z = s.reshape(self.nb, op["classes"], 128 // op["classes"]).sum(2) / (i["H"] * i["W"] * (128 // op["classes"]))
        return torch.log_softmax(z, dim=1)

    def run(self, x):
        outs = {}
        for op in self.d["program"]:
            if op["op"] == "stem":
                self.stem(op, x)
            elif op["op"] == "conv":
                self.conv(op)
            elif op["op"] == "pool":
                self.pool(op)
            else:
                outs[op["out"]] = self.head(op)
        return tuple(outs[i] for i in sorted(outs))
