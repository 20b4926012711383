"""Dependency-free PNG charts (stdlib zlib + struct) so that ``plot.py`` can always emit ``loss.png`` and
``average_elapsed_time.png`` — the two figures of the reference (plot.py:25,38) — even where matplotlib is not installed
(the B200 image).  A tiny RGB canvas with lines, rectangles and a 5x7 bitmap font for the handful of characters the two
charts need."""
from __future__ import annotations

import struct
import zlib

# 5x7 glyphs, one string of 7 rows per character ('#' = ink)
_GLYPHS = {
    "0": ["#####", "#...#", "#...#", "#...#", "#...#", "#...#", "#####"],
    "1": ["..#..", ".##..", "..#..", "..#..", "..#..", "..#..", ".###."],
    "2": ["#####", "....#", "....#", "#####", "#....", "#....", "#####"],
    "3": ["#####", "....#", "....#", ".####", "....#", "....#", "#####"],
    "4": ["#...#", "#...#", "#...#", "#####", "....#", "....#", "....#"],
    "5": ["#####", "#....", "#....", "#####", "....#", "....#", "#####"],
    "6": ["#####", "#....", "#....", "#####", "#...#", "#...#", "#####"],
    "7": ["#####", "....#", "...#.", "..#..", "..#..", "..#..", "..#.."],
    "8": ["#####", "#...#", "#...#", "#####", "#...#", "#...#", "#####"],
    "9": ["#####", "#...#", "#...#", "#####", "....#", "....#", "#####"],
    ".": [".....", ".....", ".....", ".....", ".....", ".##..", ".##.."],
    "-": [".....", ".....", ".....", "#####", ".....", ".....", "....."],
    "_": [".....", ".....", ".....", ".....", ".....", ".....", "#####"],
    " ": ["....."] * 7,
    "d": ["....#", "....#", ".####", "#...#", "#...#", "#...#", ".####"],
    "p": [".....", "####.", "#...#", "#...#", "####.", "#....", "#...."],
    "t": [".#...", ".#...", "####.", ".#...", ".#...", ".#..#", "..##."],
    "l": [".##..", "..#..", "..#..", "..#..", "..#..", "..#..", ".###."],
    "o": [".....", ".....", ".###.", "#...#", "#...#", "#...#", ".###."],
    "s": [".....", ".....", ".####", "#....", ".###.", "....#", "####."],
    "e": [".....", ".....", ".###.", "#...#", "#####", "#....", ".###."],
    "m": [".....", ".....", "##.#.", "#.#.#", "#.#.#", "#.#.#", "#.#.#"],
    "i": ["..#..", ".....", ".##..", "..#..", "..#..", "..#..", ".###."],
    "a": [".....", ".....", ".###.", "....#", ".####", "#...#", ".####"],
    "v": [".....", ".....", "#...#", "#...#", "#...#", ".#.#.", "..#.."],
    "g": [".....", ".####", "#...#", "#...#", ".####", "....#", ".###."],
    "r": [".....", ".....", "#.##.", "##..#", "#....", "#....", "#...."],
    "h": ["#....", "#....", "#.##.", "##..#", "#...#", "#...#", "#...#"],
    "c": [".....", ".....", ".###.", "#....", "#....", "#...#", ".###."],
    "[": [".###.", ".#...", ".#...", ".#...", ".#...", ".#...", ".###."],
    "]": [".###.", "...#.", "...#.", "...#.", "...#.", "...#.", ".###."],
    "/": ["....#", "....#", "...#.", "..#..", ".#...", "#....", "#...."],
}


class Canvas:
    def __init__(self, width: int, height: int, background=(255, 255, 255)):
        self.w, self.h = width, height
        self.px = bytearray(bytes(background) * (width * height))

    def set(self, x: int, y: int, color) -> None:
        if 0 <= x < self.w and 0 <= y < self.h:
            i = (y * self.w + x) * 3
            self.px[i:i + 3] = bytes(color)

    def blend(self, x: int, y: int, color, alpha: float) -> None:
        if 0 <= x < self.w and 0 <= y < self.h:
            i = (y * self.w + x) * 3
            for c in range(3):
                self.px[i + c] = int(self.px[i + c] * (1 - alpha) + color[c] * alpha)

    def line(self, x0, y0, x1, y1, color, alpha: float = 1.0) -> None:
        x0, y0, x1, y1 = int(round(x0)), int(round(y0)), int(round(x1)), int(round(y1))
        dx, dy = abs(x1 - x0), -abs(y1 - y0)
        sx, sy = (1 if x0 < x1 else -1), (1 if y0 < y1 else -1)
        err = dx + dy
        while True:
            self.blend(x0, y0, color, alpha)
            if x0 == x1 and y0 == y1:
                break
            e2 = 2 * err
            if e2 >= dy:
                err += dy
                x0 += sx
            if e2 <= dx:
                err += dx
                y0 += sy

    def rect(self, x0, y0, x1, y1, color) -> None:
        for y in range(max(0, int(min(y0, y1))), min(self.h, int(max(y0, y1)) + 1)):
            for x in range(max(0, int(min(x0, x1))), min(self.w, int(max(x0, x1)) + 1)):
                self.set(x, y, color)

    def text(self, x: int, y: int, s: str, color=(0, 0, 0), scale: int = 2) -> None:
        for ch in s:
            g = _GLYPHS.get(ch.lower(), _GLYPHS[" "])
            for r, row in enumerate(g):
                for c, bit in enumerate(row):
                    if bit == "#":
                        self.rect(x + c * scale, y + r * scale, x + c * scale + scale - 1, y + r * scale + scale - 1, color)
            x += 6 * scale

    def save(self, path: str) -> None:
        raw = b"".join(b"\x00" + bytes(self.px[y * self.w * 3:(y + 1) * self.w * 3]) for y in range(self.h))

        def chunk(tag: bytes, data: bytes) -> bytes:
            return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

        with open(path, "wb") as f:
            f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", self.w, self.h, 8, 2, 0, 0, 0)) +
                    chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def _axes(cv: Canvas, x0, y0, x1, y1, xlabel: str, ylabel: str) -> None:
    cv.line(x0, y1, x1, y1, (0, 0, 0))
    cv.line(x0, y0, x0, y1, (0, 0, 0))
    cv.text((x0 + x1) // 2 - 3 * len(xlabel) * 2, y1 + 28, xlabel)
    cv.text(4, y0 - 22, ylabel)


def line_chart(path: str, series: dict, colors: dict, xlabel: str, ylabel: str, width=800, height=500) -> None:
    """series: name -> (xs, ys).  Overlaid semi-transparent curves, min/max tick labels, colour legend."""
    cv = Canvas(width, height)
    x0, y0, x1, y1 = 70, 30, width - 20, height - 60
    xs_all = [x for xs, _ in series.values() for x in xs]
    ys_all = [y for _, ys in series.values() for y in ys]
    xmin, xmax, ymin, ymax = min(xs_all), max(xs_all), min(ys_all), max(ys_all)
    xr, yr = (xmax - xmin) or 1.0, (ymax - ymin) or 1.0
    _axes(cv, x0, y0, x1, y1, xlabel, ylabel)
    for k, frac in enumerate((0.0, 0.5, 1.0)):
        yv = ymin + frac * yr
        yy = y1 - frac * (y1 - y0)
        cv.line(x0 - 4, yy, x0, yy, (0, 0, 0))
        cv.text(4, int(yy) - 7, f"{yv:.2f}")
        xv = xmin + frac * xr
        xx = x0 + frac * (x1 - x0)
        cv.line(xx, y1, xx, y1 + 4, (0, 0, 0))
        label = f"{xv:.0f}"
        cv.text(min(int(xx) - 6 * len(label), width - 12 * len(label) - 2), y1 + 8, label)
    for i, (name, (xs, ys)) in enumerate(series.items()):
        col = colors[name]
        pts = [(x0 + (x - xmin) / xr * (x1 - x0), y1 - (y - ymin) / yr * (y1 - y0)) for x, y in zip(xs, ys)]
        for (ax, ay), (bx, by) in zip(pts, pts[1:]):
            cv.line(ax, ay, bx, by, col, alpha=0.5)
        cv.rect(x1 - 110, y0 + 10 + 22 * i, x1 - 96, y0 + 22 + 22 * i, col)
        cv.text(x1 - 90, y0 + 10 + 22 * i, name)
    cv.save(path)


def bar_chart(path: str, values: dict, colors: dict, xlabel: str, ylabel: str, width=800, height=500) -> None:
    cv = Canvas(width, height)
    x0, y0, x1, y1 = 70, 30, width - 20, height - 60
    _axes(cv, x0, y0, x1, y1, xlabel, ylabel)
    vmax = max(values.values()) or 1.0
    n = len(values)
    slot = (x1 - x0) / n
    for i, (name, v) in enumerate(values.items()):
        bx0, bx1 = x0 + slot * (i + 0.2), x0 + slot * (i + 0.8)
        by = y1 - v / vmax * (y1 - y0 - 20)
        cv.rect(bx0, by, bx1, y1 - 1, colors[name])
        cv.text(int((bx0 + bx1) / 2) - 6 * len(name), y1 + 8, name)
        cv.text(int(bx0), int(by) - 18, f"{v:.4f}")
    cv.save(path)
