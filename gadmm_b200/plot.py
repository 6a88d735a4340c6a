"""Dependency-free SVG writer for the reference's figures (``semilogy`` panels of the objective gap
against iterations, communication and clock time -- LinearRegression_Synthetic.m:146-235).

No plotting package and no graphics stack is needed: a figure is a text file.  The panels hold
exactly the arrays the drivers return (``loss_*``, ``cumulative_com_*``, ``*_time*``).
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

_COLOURS = ["#1f77b4", "#d62728", "#2ca02c", "#ff7f0e", "#9467bd", "#8c564b", "#e377c2", "#7f7f7f",
            "#bcbd22", "#17becf"]
_DASHES = ["", "6,3", "2,2", "8,3,2,3"]

Series = Dict[str, Tuple[Sequence[float], Sequence[float]]]


def _decimate(x: np.ndarray, y: np.ndarray, max_points: int = 600):
    if x.size <= max_points:
        return x, y
    idx = np.unique(np.round(np.geomspace(1, x.size, max_points)).astype(int) - 1)
    return x[idx], y[idx]


def _panel(series: Series, x0: float, y0: float, w: float, h: float, xlabel: str, ylabel: str,
           title: str, logx: bool) -> List[str]:
    xs, ys = [], []
    clean = {}
    for name, (x, y) in series.items():
        x = np.asarray(x, dtype=np.float64).reshape(-1)
        y = np.asarray(y, dtype=np.float64).reshape(-1)
        n = min(x.size, y.size)
        x, y = x[:n], y[:n]
        ok = np.isfinite(x) & np.isfinite(y) & (y > 0) & ((x > 0) if logx else True)
        if ok.sum() == 0:
            continue
        x, y = _decimate(x[ok], y[ok])
        clean[name] = (x, y)
        xs.append(x)
        ys.append(y)
    out = [f'<rect x="{x0:.1f}" y="{y0:.1f}" width="{w:.1f}" height="{h:.1f}" fill="white" stroke="black"/>',
           f'<text x="{x0 + w / 2:.1f}" y="{y0 - 8:.1f}" text-anchor="middle" font-size="13">{title}</text>',
           f'<text x="{x0 + w / 2:.1f}" y="{y0 + h + 32:.1f}" text-anchor="middle" font-size="12">{xlabel}</text>',
           f'<text transform="translate({x0 - 42:.1f},{y0 + h / 2:.1f}) rotate(-90)" text-anchor="middle" '
           f'font-size="12">{ylabel}</text>']
    if not clean:
        return out
    allx, ally = np.concatenate(xs), np.concatenate(ys)
    fx = (lambda v: np.log10(v)) if logx else (lambda v: v)
    xlo, xhi = float(fx(allx.min())), float(fx(allx.max()))
    ylo, yhi = math.floor(math.log10(ally.min())), math.ceil(math.log10(ally.max()))
    if xhi <= xlo:
        xhi = xlo + 1.0
    if yhi <= ylo:
        yhi = ylo + 1

    def px(v):
        return x0 + (fx(v) - xlo) / (xhi - xlo) * w

    def py(v):
        return y0 + h - (np.log10(v) - ylo) / (yhi - ylo) * h

    step = max(1, (yhi - ylo) // 8)
    for e in range(ylo, yhi + 1, step):
        yy = y0 + h - (e - ylo) / (yhi - ylo) * h
        out.append(f'<line x1="{x0:.1f}" y1="{yy:.1f}" x2="{x0 + w:.1f}" y2="{yy:.1f}" stroke="#dddddd"/>')
        out.append(f'<text x="{x0 - 5:.1f}" y="{yy + 4:.1f}" text-anchor="end" font-size="10">1e{e}</text>')
    for k in range(5):
        v = xlo + (xhi - xlo) * k / 4
        xx = x0 + w * k / 4
        label = f"{10 ** v:.3g}" if logx else f"{v:.3g}"
        out.append(f'<line x1="{xx:.1f}" y1="{y0 + h:.1f}" x2="{xx:.1f}" y2="{y0 + h + 4:.1f}" stroke="black"/>')
        out.append(f'<text x="{xx:.1f}" y="{y0 + h + 16:.1f}" text-anchor="middle" font-size="10">{label}</text>')
    for i, (name, (x, y)) in enumerate(clean.items()):
        pts = " ".join(f"{a:.1f},{b:.1f}" for a, b in zip(px(x), py(y)))
        dash = _DASHES[(i // len(_COLOURS)) % len(_DASHES)]
        da = f' stroke-dasharray="{dash}"' if dash else ""
        col = _COLOURS[i % len(_COLOURS)]
        out.append(f'<polyline fill="none" stroke="{col}" stroke-width="1.6"{da} points="{pts}"/>')
        ly = y0 + 14 + 14 * i
        out.append(f'<line x1="{x0 + w - 150:.1f}" y1="{ly - 4:.1f}" x2="{x0 + w - 128:.1f}" y2="{ly - 4:.1f}" '
                   f'stroke="{col}" stroke-width="1.6"{da}/>')
        out.append(f'<text x="{x0 + w - 124:.1f}" y="{ly:.1f}" font-size="10">{name}</text>')
    return out


def save_semilogy_svg(path: str, panels: Sequence[dict], width: int = 460, height: int = 340) -> str:
    """Write one SVG with ``len(panels)`` side-by-side ``semilogy`` panels.  A panel is
    ``{"series": {name: (x, y)}, "xlabel": ..., "ylabel": ..., "title": ..., "logx": bool}``."""
    n = max(len(panels), 1)
    W, H = n * width + 30, height + 30
    body = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{W}" height="{H}" viewBox="0 0 {W} {H}" '
            f'font-family="Helvetica,Arial,sans-serif">', f'<rect width="{W}" height="{H}" fill="white"/>']
    for k, p in enumerate(panels):
        body += _panel(p["series"], 70 + k * width, 30, width - 100, height - 80, p.get("xlabel", ""),
                       p.get("ylabel", "objective error"), p.get("title", ""), bool(p.get("logx", False)))
    body.append("</svg>")
    with open(path, "w") as f:
        f.write("\n".join(body))
    return path


def figure_from_curves(curves: Dict[str, np.ndarray], path: str, title: str = "") -> Optional[str]:
    """The reference's three-panel figure from a driver's curve dictionary: objective gap against
    iterations, cumulative communication (``cumulative_com_<tag>``) and time (``time_<tag>``,
    ``<tag>_time``, ``gadmm_time_<suffix>`` -- whichever the driver produced)."""
    it_s, com_s, time_s = {}, {}, {}
    for key, loss in curves.items():
        if not key.startswith("loss_"):
            continue
        tag = key[len("loss_"):]
        loss = np.asarray(loss, dtype=np.float64).reshape(-1)
        if loss.size == 0:
            continue
        it_s[tag] = (np.arange(1, loss.size + 1), loss)
        com = curves.get(f"cumulative_com_{tag}")
        if com is not None:
            com_s[tag] = (np.asarray(com).reshape(-1), loss)
        for tk in (f"time_{tag}", f"{tag}_time", "time_" + tag.replace("LAG_PS", "lagPS").replace("LAG_WK", "lagWK"),
                   "gadmm_time_" + tag.split("_", 1)[-1]):
            if tk in curves and np.asarray(curves[tk]).size:
                time_s[tag] = (np.asarray(curves[tk]).reshape(-1), loss)
                break
    if not it_s:
        return None
    panels = [{"series": it_s, "xlabel": "number of iterations", "title": title, "logx": True}]
    if com_s:
        panels.append({"series": com_s, "xlabel": "cumulative communication cost", "logx": True})
    if time_s:
        panels.append({"series": time_s, "xlabel": "clock time (s)", "logx": False})
    return save_semilogy_svg(path, panels)
