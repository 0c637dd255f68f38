"""Bar and line charts with error bars as plain SVG text (no plotting library on the bench box).

What the reference draws with matplotlib: the execution-time bar chart of ``plot()``
(``R/03 模型并行/03_model_parallel.ipynb#cell8``, json L401-431) and the split-size trade-off curve
(``#cell13``, json L599-624, artefact ``split_size_tradeoff.png``).
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence, Tuple
from xml.sax.saxutils import escape

_COLORS = ["#1f77b4", "#ff7f0e", "#2ca02c", "#d62728", "#9467bd", "#8c564b"]
W, H = 720, 440
ML, MR, MT, MB = 70, 20, 40, 70          # margins


def _nice_ticks(lo: float, hi: float, n: int = 6) -> List[float]:
    if hi <= lo:
        hi = lo + 1.0
    raw = (hi - lo) / max(n, 1)
    mag = 10 ** math.floor(math.log10(raw))
    step = min((m * mag for m in (1, 2, 2.5, 5, 10) if m * mag >= raw), default=raw)
    first = math.ceil(lo / step) * step
    out, v = [], first
    while v <= hi + 1e-12:
        out.append(round(v, 12))
        v += step
    return out


def _fmt(v: float) -> str:
    return f"{v:.3g}" if abs(v) < 1000 else f"{v:.0f}"


class _Canvas:
    def __init__(self, title: str, ylabel: str, xlabel: str, ymax: float) -> None:
        self.parts: List[str] = [
            f'<svg xmlns="http://www.w3.org/2000/svg" width="{W}" height="{H}" viewBox="0 0 {W} {H}" '
            f'font-family="sans-serif" font-size="12">',
            f'<rect width="{W}" height="{H}" fill="white"/>',
            f'<text x="{W / 2}" y="22" text-anchor="middle" font-size="15">{escape(title)}</text>',
            f'<text x="16" y="{(MT + H - MB) / 2}" text-anchor="middle" transform="rotate(-90 16 {(MT + H - MB) / 2})">'
            f'{escape(ylabel)}</text>',
            f'<text x="{(ML + W - MR) / 2}" y="{H - 12}" text-anchor="middle">{escape(xlabel)}</text>']
        self.ymax = ymax if ymax > 0 else 1.0
        for t in _nice_ticks(0.0, self.ymax):
            y = self.y(t)
            self.parts.append(f'<line x1="{ML}" y1="{y:.1f}" x2="{W - MR}" y2="{y:.1f}" stroke="#dddddd"/>')
            self.parts.append(f'<text x="{ML - 6}" y="{y + 4:.1f}" text-anchor="end">{_fmt(t)}</text>')
        self.parts.append(f'<line x1="{ML}" y1="{MT}" x2="{ML}" y2="{H - MB}" stroke="black"/>')
        self.parts.append(f'<line x1="{ML}" y1="{H - MB}" x2="{W - MR}" y2="{H - MB}" stroke="black"/>')

    def y(self, v: float) -> float:
        return H - MB - (H - MB - MT) * (v / self.ymax)

    def errbar(self, x: float, mean: float, std: float, color: str = "black") -> None:
        if std <= 0:
            return
        y0, y1 = self.y(max(mean - std, 0.0)), self.y(mean + std)
        self.parts.append(f'<line x1="{x:.1f}" y1="{y0:.1f}" x2="{x:.1f}" y2="{y1:.1f}" stroke="{color}"/>')
        for yy in (y0, y1):
            self.parts.append(f'<line x1="{x - 4:.1f}" y1="{yy:.1f}" x2="{x + 4:.1f}" y2="{yy:.1f}" stroke="{color}"/>')

    def legend(self, names: Sequence[str]) -> None:
        for i, nm in enumerate(names):
            y = MT + 8 + 16 * i
            self.parts.append(f'<rect x="{W - MR - 200}" y="{y - 9}" width="12" height="10" fill="{_COLORS[i % len(_COLORS)]}"/>')
            self.parts.append(f'<text x="{W - MR - 183}" y="{y}">{escape(nm)}</text>')

    def done(self) -> str:
        return "\n".join(self.parts + ["</svg>"]) + "\n"


def bar_chart(means: Sequence[float], stds: Sequence[float], labels: Sequence[str], title: str = "",
              ylabel: str = "", path: Optional[str] = None) -> str:
    """One bar per label with a +-std error bar (the reference's ``plot(means, stds, labels, fig_name)``)."""
    assert len(means) == len(stds) == len(labels) and len(means) > 0
    c = _Canvas(title, ylabel, "", max(m + s for m, s in zip(means, stds)) * 1.1)
    slot = (W - ML - MR) / len(means)
    for i, (m, s, lb) in enumerate(zip(means, stds, labels)):
        x0, bw = ML + slot * i + slot * 0.2, slot * 0.6
        c.parts.append(f'<rect x="{x0:.1f}" y="{c.y(m):.1f}" width="{bw:.1f}" height="{c.y(0) - c.y(m):.1f}" '
                       f'fill="{_COLORS[i % len(_COLORS)]}" fill-opacity="0.6"/>')
        c.errbar(x0 + bw / 2, m, s)
        c.parts.append(f'<text x="{x0 + bw / 2:.1f}" y="{c.y(m + s) - 5:.1f}" text-anchor="middle">{_fmt(m)}</text>')
        c.parts.append(f'<text x="{x0 + bw / 2:.1f}" y="{H - MB + 16}" text-anchor="middle">{escape(lb)}</text>')
    svg = c.done()
    if path:
        with open(path, "w") as f:
            f.write(svg)
    return svg


def line_chart(series: Sequence[Tuple[str, Sequence[float], Sequence[float], Sequence[float]]], title: str = "",
               xlabel: str = "", ylabel: str = "", path: Optional[str] = None) -> str:
    """``series`` = (name, xs, means, stds) each; xs are placed at equal distance (like the reference's
    ``range(len(split_sizes))`` with the split sizes as tick labels)."""
    assert series and all(len(xs) == len(ms) == len(ss) and len(xs) > 0 for _, xs, ms, ss in series)
    ticks = list(series[0][1])
    c = _Canvas(title, ylabel, xlabel, max(m + s for _, _, ms, ss in series for m, s in zip(ms, ss)) * 1.1)
    n = len(ticks)
    px = lambda i: ML + (W - ML - MR) * ((i + 0.5) / n)  # noqa: E731
    for i, t in enumerate(ticks):
        c.parts.append(f'<text x="{px(i):.1f}" y="{H - MB + 16}" text-anchor="middle">{escape(str(t))}</text>')
    for k, (_, xs, ms, ss) in enumerate(series):
        col = _COLORS[k % len(_COLORS)]
        idx = [ticks.index(x) for x in xs]
        pts = " ".join(f"{px(i):.1f},{c.y(m):.1f}" for i, m in zip(idx, ms))
        c.parts.append(f'<polyline points="{pts}" fill="none" stroke="{col}" stroke-width="2"/>')
        for i, m, s in zip(idx, ms, ss):
            c.errbar(px(i), m, s, col)
            c.parts.append(f'<circle cx="{px(i):.1f}" cy="{c.y(m):.1f}" r="3.5" fill="{col}"/>')
    c.legend([nm for nm, _, _, _ in series])
    svg = c.done()
    if path:
        with open(path, "w") as f:
            f.write(svg)
    return svg
