"""Oracle (test infrastructure): the ColorModel-shaped layer graph on the CPU, in canonical row order.

Follows the call surface the reference uses: g_a(x) -> (y, k)
(codec_pipeline.py:273), entropy_model.h_a / h_s (:287, :354), g_s.down_conv (codec_parallel.py:302-303),
g_s(y_hat, k=ks) (:469).  The layer graph itself is this repository's canonical stand-in for the missing
`demo_small` model; weights come in as a state dict of numpy arrays.

`bf16=True` emulates what the tensor-core path computes: conv inputs and weights rounded to bfloat16,
fp32 accumulation, activations stored as bfloat16 between layers; layer outputs the product keeps in
fp32 (y, z, Gaussian params, occupancy logits, colours) are left in fp32.
"""
import numpy as np

from . import sparse_ref as S


class RefModel:
    def __init__(self, cfg, sd, bf16=True):
        self.cfg = cfg
        self.sd = sd
        self.bf16 = bf16
        self._kmaps = {}

    # ---- helpers -------------------------------------------------------------------------------------
    def _store(self, F):
        return S.bf16_round(F) if self.bf16 else F

    def _conv(self, name, F, nbr, relu=False, residual=None, store=True):
        out = S.conv_forward(F, nbr, self.sd[name + ".kernel"], self.sd[name + ".bias"], residual=residual,
                             relu=relu, bf16=self.bf16)
        return self._store(out) if store else out

    def _k3(self, C, stride):
        key = (C.shape[0], int(stride), C[:4].tobytes(), C[-4:].tobytes())
        if key not in self._kmaps:
            self._kmaps[key] = S.kernel_map(C, C, S.region_offsets(3, stride))
        return self._kmaps[key]

    def _down(self, name, C, F, stride, relu, store=True):
        """k=2, s=2 strided conv from tensor stride `stride` to 2*stride."""
        Co = S.stride_coords(C, 2 * stride)
        nbr = S.kernel_map(C, Co, S.region_offsets(2, stride))
        return Co, self._conv(name, F, nbr, relu=relu, store=store)

    def _up(self, name, C, F, stride, relu):
        """k=2, s=2 generative transposed conv from tensor stride `stride` to stride/2."""
        t = stride // 2
        Co = S.generative_coords(C, t)
        nbr = S.kernel_map(C, Co, S.region_offsets(2, t), transposed=True)
        return Co, self._conv(name, F, nbr, relu=relu)

    def _block(self, prefix, C, F, stride):
        nbr = self._k3(C, stride)
        h = self._conv(prefix + ".conv0", F, nbr, relu=True)
        return self._conv(prefix + ".conv1", h, nbr, relu=True, residual=F)

    @staticmethod
    def frame_counts(C, num_frames):
        return [int((C[:, 0] == b).sum()) for b in range(num_frames)]

    # ---- transforms ------------------------------------------------------------------------------------
    def g_a(self, C, F, num_frames):
        """-> (C_y, y [Ny, Cy] fp32, k) with k[level][frame] = rows to keep after up-stage `level`."""
        cfg = self.cfg
        F = self._store(np.asarray(F, np.float32))
        F = self._conv("g_a.stem", F, self._k3(C, 1), relu=True)
        counts = []
        stride = 1
        for i in range(len(cfg["widths"])):
            counts.append(self.frame_counts(C, num_frames))
            C, F = self._down(f"g_a.down{i}", C, F, stride, relu=True)
            stride *= 2
            for b in range(cfg["blocks"]):
                F = self._block(f"g_a.block{i}_{b}", C, F, stride)
        y = self._conv("g_a.head", F, None, store=False)
        return C, y, counts[::-1]

    def h_a(self, C, y):
        F = self._store(np.asarray(y, np.float32))
        C, F = self._down("entropy_model.h_a.down0", C, F, 8, relu=True)
        C, z = self._down("entropy_model.h_a.down1", C, F, 16, relu=False, store=False)
        return C, z

    def down_conv_coords(self, C, stride):
        return S.stride_coords(C, 2 * stride)

    def h_s(self, C, z_hat):
        F = self._store(np.asarray(z_hat, np.float32))
        C, F = self._up("entropy_model.h_s.up0", C, F, 32, relu=True)
        C, F = self._up("entropy_model.h_s.up1", C, F, 16, relu=True)
        params = self._conv("entropy_model.h_s.head", F, None, store=False)
        return C, params

    def g_s_pre(self, y_hat):
        """g_s's 1x1 input convolution on y_hat [Ny, Cy] (rows in any order: no neighbours involved)."""
        return self._conv("g_s.pre", self._store(np.asarray(y_hat, np.float32)), None, relu=True)

    def g_s_stage(self, i, C, F, stride):
        """Up-stage i of g_s on GIVEN inputs (coordinates at tensor stride `stride`, their features):
        generative up-convolution, residual blocks, occupancy logits on the un-pruned candidates.
        -> (C_children at stride/2, F [n, c], logits [n, 1]).  Stage-wise parity tests feed the product's
        kept rows of stage i-1 here (codec_parallel.py:469 runs all three inside g_s)."""
        C, F = self._up(f"g_s.up{i}", C, F, stride, relu=True)
        stride //= 2
        for b in range(self.cfg["blocks"]):
            F = self._block(f"g_s.block{i}_{b}", C, F, stride)
        logits = self._conv(f"g_s.occ{i}", F, self._k3(C, stride), store=False)
        return C, F, logits

    def g_s_color(self, C, F):
        """Colour head on the final (stride-1) coordinates: fp32 [n, 3], before pack_batches' clip."""
        return self._conv("g_s.color", F, self._k3(C, 1), store=False)

    def g_s(self, C, y_hat, ks, return_intermediates=False, keeps=None):
        """keeps: optional list of three boolean masks (canonical row order of each level's candidates) that REPLACE
        the oracle's own top-k decisions -- parity tests pass the product's decisions so that both sides reconstruct
        the same geometry and the attributes can be compared point by point."""
        F = self.g_s_pre(y_hat)
        stride = 8
        inter = []
        for i in range(3):
            C, F, logits = self.g_s_stage(i, C, F, stride)
            stride //= 2
            keep = S.topk_mask(logits[:, 0], C, ks[i]) if keeps is None else np.asarray(keeps[i], bool)
            if return_intermediates:
                inter.append({"C": C, "F": F, "logits": logits, "keep": keep})
            C, F = C[keep], F[keep]
        colors = self.g_s_color(C, F)
        if return_intermediates:
            return C, colors, inter
        return C, colors
