"""Digitise the reference's own result figures into golden curves (TEST INFRASTRUCTURE ONLY).

The reference stores no numbers, but its committed matplotlib figures are vector plots whose
polyline vertices are plain numbers once the PDF content stream is inflated:

  fig/fig_prox_lmc_deconv_einstein_{MYULA,ULPDA,MAP}_1000_snr_psnr_mse.pdf
      written by /root/reference/prox_lmc_deconv.py:810-853 (samplers) and :402-445 (MAP):
      three stacked axes (SNR, PSNR, MSE of every *sample* / MAP iterate against the ground truth),
      nine curves each (models M1..M9 in the order of lines 813-821), x = sample index 0..N-1.
      The run is fixed by /root/reference/expt_deconv.sh:4-6: --image=einstein --sigma=0.75
      --tau=0.3 --gamma_mc=15 --gamma_me=15 --N=1000 (seed 0, niter_l2=50, niter_tv=10).
  fig/fig_n5_gamma0.1_10000_wass_dist.pdf etc.
      written by /root/reference/lmc.py:418-443 / lmc_laplace.py:453-476: W2 distance of chain
      prefixes against true samples, one curve per sampler.

matplotlib's path simplifier keeps only the vertices needed to draw the curve within 1/9 px, so a
1000-sample curve survives as 15-35 vertices; every kept vertex is an original data point
(x lands on an integer sample index to ~1e-4), and the PDF quotes coordinates to 1e-6 pt, i.e.
about 2e-7 dB on the SNR/PSNR axes.  The axes are calibrated from the tick labels (real text in
the PDF) and the tick marks drawn next to them.

Usage (in the build container, where /root/reference exists):
    python oracle/digitise_figs.py            # writes tests/golden/deconv_curves.npz, w2_curves.npz
Nothing on the GPU box reads /root/reference; the committed .npz files are the fixtures.
"""
import os
import re
import sys
import zlib

import numpy as np

REF = os.environ.get("LMC_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(os.path.dirname(HERE), "tests", "golden")


def page_stream(path):
    """Inflate the page content stream (the first stream that draws lines)."""
    d = open(path, "rb").read()
    for m in re.finditer(rb"stream\r?\n", d):
        s = m.end()
        e = d.find(b"endstream", s)
        try:
            txt = zlib.decompress(d[s:e])
        except zlib.error:
            continue
        if b" l\n" in txt and b" re W n" in txt:
            return txt.decode("latin1")
    raise RuntimeError(f"no drawable stream in {path}")


def _isnum(t):
    try:
        float(t)
        return True
    except ValueError:
        return False


def parse(stream):
    """Walk the PDF operators; return (curves, ticks, labels).

    curves: list of dict(clip=(x,y,w,h), rgb, dash, pts (n,2)) for stroked paths with > 2 vertices
            inside an axes clip rectangle;
    ticks : list of ((x0,y0),(x1,y1)) two-point segments painted with `B` (tick marks);
    labels: list of (x, y, text) from `Td`/`Tm` + `Tj`.
    """
    toks = re.findall(r"\((?:[^()\\]|\\.)*\)|\[|\]|[^\s\[\]()]+", stream)
    stack, cur, curves, ticks, labels = [], [], [], [], []
    gs = {"clip": None, "rgb": (0.0, 0.0, 0.0), "dash": ()}
    saved = []
    pend_rect, text_pos, in_arr, arr = None, None, False, []
    for t in toks:
        if t == "[":
            in_arr, arr = True, []
            continue
        if t == "]":
            in_arr = False
            stack.append(tuple(arr))
            continue
        if in_arr:
            arr.append(float(t))
            continue
        if _isnum(t):
            stack.append(float(t))
            continue
        if t.startswith("(") or t.startswith("/"):
            stack.append(t)
            continue
        op = t
        if op == "q":
            saved.append(dict(gs))
        elif op == "Q":
            gs = saved.pop() if saved else gs
        elif op == "re":
            pend_rect = tuple(stack[-4:])
        elif op == "W":
            gs["clip"] = pend_rect
        elif op == "RG":
            gs["rgb"] = tuple(stack[-3:])
        elif op == "G":
            gs["rgb"] = (stack[-1],) * 3
        elif op == "d":
            gs["dash"] = stack[-2] if isinstance(stack[-2], tuple) else ()
        elif op == "m":
            cur = [tuple(stack[-2:])]
        elif op == "l":
            cur.append(tuple(stack[-2:]))
        elif op == "S":
            if len(cur) > 2 and gs["clip"] is not None:
                pts = np.array(cur)
                if len(pts) > 1 and np.all(pts[-1] == pts[-2]):
                    pts = pts[:-1]          # the pdf backend repeats the last vertex
                curves.append(dict(clip=gs["clip"], rgb=gs["rgb"], dash=gs["dash"], pts=pts))
            cur = []
        elif op == "B":
            if len(cur) == 2:
                ticks.append((cur[0], cur[1]))
            cur = []
        elif op in ("f", "n", "h"):
            if op != "h":
                cur = []
        elif op == "Td":
            if text_pos is None:
                text_pos = tuple(stack[-2:])
            elif stack[-2:] != [0.0, 0.0]:
                text_pos = tuple(stack[-2:])
        elif op == "Tm":
            text_pos = tuple(stack[-2:])
        elif op == "BT":
            text_pos = None
        elif op == "Tj":
            s = stack[-1]
            labels.append((text_pos[0], text_pos[1], s[1:-1]))
        if op not in ("re",):
            stack = []
    return curves, ticks, labels


def calibrate_axes(clip, ticks, labels):
    """Linear data<-page maps of one axes from its major tick marks and numeric tick labels."""
    x0, y0, w, h = clip
    eps = 1e-3

    def numeric(s):
        s = s.replace("\\(", "(").replace("\\)", ")")
        # usetex minus sign arrives as a non-ASCII glyph; accept ASCII '-' and the cmsy glyph \000/\241
        s2 = re.sub(r"^(\\[0-7]{3}|[^0-9.])+", "-", s) if not s[0].isdigit() else s
        try:
            return float(s2)
        except ValueError:
            return None

    # major ticks have length 3 pt, minor 1.5 pt (scienceplots style)
    ymaj = sorted({round(a[1], 6) for a, b in ticks
                   if abs(a[0] - x0) < eps and abs(a[1] - b[1]) < eps and abs(abs(b[0] - a[0]) - 3.0) < 0.01})
    xmaj = sorted({round(a[0], 6) for a, b in ticks
                   if abs(a[1] - y0) < eps and abs(a[0] - b[0]) < eps and abs(abs(b[1] - a[1]) - 3.0) < 0.01})
    ylab = [(ly, numeric(s)) for lx, ly, s in labels
            if lx < x0 and y0 - 5 < ly < y0 + h + 5 and numeric(s) is not None]
    xlab = [(lx, numeric(s)) for lx, ly, s in labels
            if y0 - 14 < ly < y0 - 6 and x0 - 10 < lx < x0 + w + 10 and numeric(s) is not None]

    def fit(major, lab, centre_guess):
        # pair each label with the nearest major tick after removing the constant text offset
        lab = sorted(lab)
        pos = np.array([p for p, v in lab])
        val = np.array([v for p, v in lab])
        best = None
        for off in centre_guess:
            idx = [int(np.argmin(np.abs(np.array(major) - (p + off)))) for p in pos]
            if len(set(idx)) != len(idx):
                continue
            tp = np.array([major[i] for i in idx])
            a, b = np.polyfit(tp, val, 1)
            res = np.max(np.abs(a * tp + b - val))
            if best is None or res < best[0]:
                best = (res, a, b)
        assert best is not None and best[0] < 1e-6 * max(1.0, np.max(np.abs(val))), (best, lab, major)
        return best[1], best[2]

    ay, by = fit(ymaj, ylab, [3.377, 3.4, 3.3])
    # x labels are centred on the tick: offset = half the label width; fit through a rough pairing
    ax, bx = fit(xmaj, xlab, [2.49, 7.47, 9.96, 4.98, 12.45])
    return (ax, bx), (ay, by)


def digitise(path, n_axes, n_curves):
    """-> list over axes (top to bottom) of list over curves of (x_data, y_data) vertex arrays."""
    curves, ticks, labels = parse(page_stream(path))
    clips = sorted({c["clip"] for c in curves}, key=lambda r: -r[1])
    assert len(clips) == n_axes, (path, clips)
    out = []
    for clip in clips:
        (ax, bx), (ay, by) = calibrate_axes(clip, ticks, labels)
        cs = [c for c in curves if c["clip"] == clip]
        assert len(cs) == n_curves, (path, clip, len(cs))
        out.append([(ax * c["pts"][:, 0] + bx, ay * c["pts"][:, 1] + by) for c in cs])
    return out


DECONV_FIGS = {
    "einstein_MYULA": "fig/fig_prox_lmc_deconv_einstein_MYULA_1000_snr_psnr_mse.pdf",
    "einstein_ULPDA": "fig/fig_prox_lmc_deconv_einstein_ULPDA_1000_snr_psnr_mse.pdf",
    "einstein_MAP": "fig/fig_prox_lmc_deconv_einstein_MAP_1000_snr_psnr_mse.pdf",
    "camera_MYULA": "fig/fig_prox_lmc_deconv_camera_MYULA_1000_snr_psnr_mse.pdf",
    "camera_ULPDA": "fig/fig_prox_lmc_deconv_camera_ULPDA_1000_snr_psnr_mse.pdf",
    "camera_MAP": "fig/fig_prox_lmc_deconv_camera_MAP_1000_snr_psnr_mse.pdf",
}
METRICS = ("snr", "psnr", "mse")


def deconv_curves():
    """{f"{fig}_{metric}_M{k}_idx"/"_val"} for every figure, metric and model."""
    out = {}
    for name, rel in DECONV_FIGS.items():
        p = os.path.join(REF, rel)
        if not os.path.exists(p):
            continue
        axes = digitise(p, 3, 9)
        for metric, curves in zip(METRICS, axes):
            for k, (x, y) in enumerate(curves, start=1):
                idx = np.rint(x).astype(np.int64)
                # every kept vertex must sit on an integer sample index
                assert np.max(np.abs(x - idx)) < 2e-3, (name, metric, k, np.max(np.abs(x - idx)))
                out[f"{name}_{metric}_M{k}_idx"] = idx
                out[f"{name}_{metric}_M{k}_val"] = y
    return out


if __name__ == "__main__":
    os.makedirs(GOLDEN, exist_ok=True)
    dc = deconv_curves()
    np.savez_compressed(os.path.join(GOLDEN, "deconv_curves.npz"), **dc)
    print("deconv_curves.npz:", len(dc) // 2, "curves")
    for fig in ("einstein_MYULA", "einstein_ULPDA", "einstein_MAP"):
        for m in METRICS:
            k = f"{fig}_{m}_M1"
            if k + "_idx" in dc:
                print(k, dc[k + "_idx"][:6], np.round(dc[k + "_val"][:6], 5), "... last",
                      dc[k + "_idx"][-1], round(float(dc[k + "_val"][-1]), 5))
