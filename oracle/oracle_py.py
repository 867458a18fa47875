"""TEST INFRASTRUCTURE — ctypes wrappers around oracle/liboracle.so (CPU restatement) and oracle/_ref/libref_rast.so
(the unmodified reference CUDA built for sm_100a).  Imported only by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs; never by rain_b200."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libref_rast.so")


def build(ref=True):
    subprocess.check_call(["make", "-s", "-f", os.path.join(HERE, "Makefile"), "oracle"])
    if ref:
        subprocess.check_call(["make", "-s", "-f", os.path.join(HERE, "Makefile"), "ref"])


def _fp(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


class Oracle:
    """CPU oracle; all arrays are numpy."""
    _names = {"depths": np.float32, "means2D": np.float32, "cov3D": np.float32, "conic_opacity": np.float32, "rgb": np.float32,
              "clamped": np.uint8, "tiles_touched": np.uint32, "point_offsets": np.uint32, "keys_unsorted": np.uint64,
              "values_unsorted": np.uint32, "keys": np.uint64, "point_list": np.uint32, "ranges": np.uint32,
              "final_T": np.float32, "n_contrib": np.uint32, "radii": np.int32}

    def __init__(self):
        if not os.path.exists(ORACLE_SO):
            build(ref=False)
        self.lib = C.CDLL(ORACLE_SO)
        L = self.lib
        L.orc_create.restype = C.c_void_p
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_forward.restype = C.c_int64
        L.orc_forward.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 5 + \
            [C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_backward.restype = C.c_int
        L.orc_backward.argtypes = [C.c_void_p] * 11
        L.orc_get.restype = C.c_int64
        L.orc_get.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_void_p)]
        L.orc_counters.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
        L.orc_mark_visible.argtypes = [C.c_int] + [C.c_void_p] * 4
        L.orc_dist2.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        L.orc_dist2_brute.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        L.orc_dist2_queries.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_morton.restype = C.c_uint32
        L.orc_morton.argtypes = [C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        L.orc_num_threads.restype = C.c_int
        self.ctx = L.orc_create()
        self._shape = None

    def __del__(self):
        try:
            self.lib.orc_destroy(self.ctx)
        except Exception:
            pass

    def num_threads(self):
        return self.lib.orc_num_threads()

    def forward(self, bg, means3D, opacities, scales, rotations, viewmatrix, projmatrix, campos, tanfovx, tanfovy, H, W,
                shs=None, sh_degree=0, colors_precomp=None, cov3D_precomp=None, scale_modifier=1.0, low_pass=0.3):
        means3D = _f32(means3D)
        P = means3D.shape[0]
        shs, colors_precomp, cov3D_precomp, scales, rotations = map(_f32, (shs, colors_precomp, cov3D_precomp, scales, rotations))
        M = shs.shape[1] if shs is not None else 0
        out_color = np.zeros((3, H, W), np.float32); out_depth = np.zeros((1, H, W), np.float32); radii = np.zeros((P,), np.int32)
        keep = [_f32(bg), means3D, shs, colors_precomp, _f32(opacities), scales, rotations, cov3D_precomp, _f32(viewmatrix), _f32(projmatrix), _f32(campos)]
        N = self.lib.orc_forward(self.ctx, P, int(sh_degree), M, _fp(keep[0]), W, H, _fp(means3D), _fp(shs), _fp(colors_precomp), _fp(keep[4]),
                                 _fp(scales), float(scale_modifier), _fp(rotations), _fp(cov3D_precomp), _fp(keep[8]), _fp(keep[9]), _fp(keep[10]),
                                 float(tanfovx), float(tanfovy), float(low_pass), _fp(out_color), _fp(out_depth), _fp(radii))
        if N < 0:
            raise RuntimeError("orc_forward failed")
        self._shape = (P, M, H, W)
        return int(N), out_color, radii, out_depth

    def backward(self, dL_dpix):
        P, M, H, W = self._shape
        dL_dpix = _f32(dL_dpix)
        z = lambda *s: np.zeros(s, np.float32)
        g = dict(dL_dmean2D=z(P, 3), dL_dconic=z(P, 4), dL_dopacity=z(P, 1), dL_dcolor=z(P, 3), dL_dmean3D=z(P, 3), dL_dcov3D=z(P, 6),
                 dL_dsh=z(P, max(M, 1), 3), dL_dscale=z(P, 3), dL_drot=z(P, 4))
        self.lib.orc_backward(self.ctx, _fp(dL_dpix), _fp(g["dL_dmean2D"]), _fp(g["dL_dconic"]), _fp(g["dL_dopacity"]), _fp(g["dL_dcolor"]),
                              _fp(g["dL_dmean3D"]), _fp(g["dL_dcov3D"]), _fp(g["dL_dsh"]), _fp(g["dL_dscale"]), _fp(g["dL_drot"]))
        if M == 0:
            g["dL_dsh"] = z(P, 0, 3)
        return g

    def get(self, name):
        p = C.c_void_p()
        n = self.lib.orc_get(self.ctx, name.encode(), C.byref(p))
        if n < 0:
            raise KeyError(name)
        dt = np.dtype(self._names[name])
        if n == 0:
            return np.zeros((0,), dt)
        buf = (C.c_char * (n * dt.itemsize)).from_address(p.value)
        return np.frombuffer(buf, dtype=dt).copy()

    def counters(self):
        out = (C.c_int64 * 4)()
        self.lib.orc_counters(self.ctx, out)
        return dict(fwd_evals=out[0], fwd_contribs=out[1], bwd_evals=out[2], bwd_contribs=out[3])

    def mark_visible(self, means3D, viewmatrix, projmatrix):
        means3D = _f32(means3D); v = _f32(viewmatrix); p = _f32(projmatrix)
        out = np.zeros((means3D.shape[0],), np.uint8)
        self.lib.orc_mark_visible(means3D.shape[0], _fp(means3D), _fp(v), _fp(p), _fp(out))
        return out.astype(bool)

    def dist2(self, points, brute=False):
        points = _f32(points)
        out = np.zeros((points.shape[0],), np.float32)
        (self.lib.orc_dist2_brute if brute else self.lib.orc_dist2)(points.shape[0], _fp(points), _fp(out))
        return out

    def dist2_queries(self, points, idx):
        points = _f32(points); idx = np.ascontiguousarray(idx, dtype=np.int32)
        out = np.zeros((idx.shape[0],), np.float32)
        self.lib.orc_dist2_queries(points.shape[0], _fp(points), idx.shape[0], _fp(idx), _fp(out))
        return out

    def morton(self, x, y, z, minn, maxx):
        a = _f32(minn); b = _f32(maxx)
        return int(self.lib.orc_morton(float(x), float(y), float(z), _fp(a), _fp(b)))


class RefCuda:
    """The unmodified reference CUDA (oracle/_ref/libref_rast.so) driven with torch CUDA tensors.  Plays the role of the
    reference's torch binding (rasterize_points.cu): allocates/zero-fills outputs exactly as it does, then calls the
    reference's Rasterizer through the shim.  All launches go to the legacy default stream, as in the reference."""

    def __init__(self):
        import torch
        self.torch = torch
        if not os.path.exists(REF_SO):
            raise RuntimeError(f"{REF_SO} missing: build it in the container with `make -f oracle/Makefile ref`")
        self.lib = C.CDLL(REF_SO)
        L = self.lib
        vp, i, f = C.c_void_p, C.c_int, C.c_float
        L.ref_forward.restype = i
        L.ref_forward.argtypes = [i, i, i, vp, i, i, vp, vp, vp, vp, vp, f, vp, vp, vp, vp, vp, f, f, i, vp, vp, f, vp, i]
        L.ref_backward.restype = i
        L.ref_backward.argtypes = [i, i, i, i, vp, i, i, vp, vp, vp, vp, f, vp, vp, vp, vp, vp, f, f, vp, vp] + [vp] * 9 + [i, f]
        L.ref_mark_visible.argtypes = [i, vp, vp, vp, vp]
        L.ref_state_ptrs.restype = i
        L.ref_state_ptrs.argtypes = [C.POINTER(vp)]
        L.ref_dist2.argtypes = [i, vp, vp]
        L.ref_last_error.restype = C.c_char_p
        self._last = None

    @staticmethod
    def _p(t):
        return None if (t is None or t.numel() == 0) else t.data_ptr()

    def forward(self, bg, means3D, opacities, scales, rotations, viewmatrix, projmatrix, campos, tanfovx, tanfovy, H, W,
                shs=None, sh_degree=0, colors_precomp=None, cov3D_precomp=None, scale_modifier=1.0, low_pass=0.3, debug=False, sync=True):
        torch = self.torch
        P = means3D.shape[0]
        dev = means3D.device
        M = shs.shape[1] if shs is not None else 0
        out_color = torch.full((3, H, W), 0.0, dtype=torch.float32, device=dev)   # rasterize_points.cu:58-60
        radii = torch.full((P,), 0, dtype=torch.int32, device=dev)
        out_depth = torch.full((1, H, W), 0.0, dtype=torch.float32, device=dev)
        if sync:    # only needed when the caller is not on torch's default stream (== the legacy stream the reference uses)
            torch.cuda.current_stream().synchronize()
        N = self.lib.ref_forward(P, int(sh_degree), M, self._p(bg), W, H, self._p(means3D), self._p(shs), self._p(colors_precomp),
                                 self._p(opacities), self._p(scales), float(scale_modifier), self._p(rotations), self._p(cov3D_precomp),
                                 self._p(viewmatrix), self._p(projmatrix), self._p(campos), float(tanfovx), float(tanfovy), 0,
                                 out_color.data_ptr(), out_depth.data_ptr(), float(low_pass), radii.data_ptr(), int(debug))
        if N < 0:
            raise RuntimeError("ref_forward: " + self.lib.ref_last_error().decode())
        self._last = dict(P=P, M=M, H=H, W=W, N=N, D=int(sh_degree), bg=bg, means3D=means3D, shs=shs, colors_precomp=colors_precomp,
                          scales=scales, rotations=rotations, cov3D_precomp=cov3D_precomp, view=viewmatrix, proj=projmatrix, campos=campos,
                          tanfovx=tanfovx, tanfovy=tanfovy, scale_modifier=scale_modifier, low_pass=low_pass, radii=radii, dev=dev)
        return int(N), out_color, radii, out_depth

    def backward(self, dL_dpix, debug=False, sync=True):
        torch = self.torch
        s = self._last
        P, M, dev = s["P"], s["M"], s["dev"]
        z = lambda *sh: torch.zeros(sh, dtype=torch.float32, device=dev)                # rasterize_points.cu:145-153
        g = dict(dL_dmean3D=z(P, 3), dL_dmean2D=z(P, 3), dL_dcolor=z(P, 3), dL_dconic=z(P, 2, 2), dL_dopacity=z(P, 1), dL_dcov3D=z(P, 6),
                 dL_dsh=z(P, M, 3), dL_dscale=z(P, 3), dL_drot=z(P, 4))
        if sync:
            torch.cuda.current_stream().synchronize()
        r = self.lib.ref_backward(P, s["D"], M, s["N"], self._p(s["bg"]), s["W"], s["H"], self._p(s["means3D"]), self._p(s["shs"]),
                                  self._p(s["colors_precomp"]), self._p(s["scales"]), float(s["scale_modifier"]), self._p(s["rotations"]),
                                  self._p(s["cov3D_precomp"]), self._p(s["view"]), self._p(s["proj"]), self._p(s["campos"]),
                                  float(s["tanfovx"]), float(s["tanfovy"]), s["radii"].data_ptr(), dL_dpix.data_ptr(),
                                  g["dL_dmean2D"].data_ptr(), g["dL_dconic"].data_ptr(), g["dL_dopacity"].data_ptr(), g["dL_dcolor"].data_ptr(),
                                  g["dL_dmean3D"].data_ptr(), g["dL_dcov3D"].data_ptr(), self._p(g["dL_dsh"]), g["dL_dscale"].data_ptr(),
                                  g["dL_drot"].data_ptr(), int(debug), float(s["low_pass"]))
        if r != 0:
            raise RuntimeError("ref_backward: " + self.lib.ref_last_error().decode())
        return g

    def state(self):
        """Copies of the reference's private scratch arrays of the last forward (torch tensors on the device)."""
        torch = self.torch
        s = self._last
        P, N, H, W, dev = s["P"], s["N"], s["H"], s["W"], s["dev"]
        T = ((W + 15) // 16) * ((H + 15) // 16)
        ptrs = (C.c_void_p * 16)()
        if self.lib.ref_state_ptrs(ptrs) != 0:
            raise RuntimeError("ref_state_ptrs failed")
        torch.cuda.synchronize()
        rt = C.CDLL("libcudart.so.12")      # already in the process (libref_rast.so links it); same soname -> same handle
        rt.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]

        def grab(i, shape, dtype):
            t = torch.empty(shape, dtype=dtype, device=dev)
            if t.numel():
                if rt.cudaMemcpy(t.data_ptr(), ptrs[i], t.numel() * t.element_size(), 3) != 0:
                    raise RuntimeError("cudaMemcpy of reference state failed")
            return t
        st = dict(depths=grab(0, (P,), torch.float32), clamped=grab(1, (P, 3), torch.uint8), means2D=grab(2, (P, 2), torch.float32),
                  cov3D=grab(3, (P, 6), torch.float32), conic_opacity=grab(4, (P, 4), torch.float32), rgb=grab(5, (P, 3), torch.float32),
                  tiles_touched=grab(6, (P,), torch.int32), point_offsets=grab(7, (P,), torch.int32),
                  final_T=grab(8, (H * W,), torch.float32), n_contrib=grab(9, (H * W,), torch.int32), ranges=grab(10, (T, 2), torch.int32))
        if N > 0:
            st["keys"] = grab(11, (N,), torch.int64)
            st["point_list"] = grab(12, (N,), torch.int32)
        torch.cuda.synchronize()
        return st

    def mark_visible(self, means3D, viewmatrix, projmatrix):
        torch = self.torch
        present = torch.zeros((means3D.shape[0],), dtype=torch.bool, device=means3D.device)
        torch.cuda.current_stream().synchronize()
        self.lib.ref_mark_visible(means3D.shape[0], means3D.data_ptr(), viewmatrix.data_ptr(), projmatrix.data_ptr(), present.data_ptr())
        torch.cuda.synchronize()
        return present

    def dist2(self, points):
        torch = self.torch
        out = torch.full((points.shape[0],), 0.0, dtype=torch.float32, device=points.device)
        torch.cuda.current_stream().synchronize()
        self.lib.ref_dist2(points.shape[0], points.data_ptr(), out.data_ptr())
        torch.cuda.synchronize()
        return out
