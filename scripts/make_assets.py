"""Writes the small LDR texture the textured scene variant maps onto surfaces: assets/data/tiles.bmp, a 64x32 24-bit
uncompressed BMP (every texel a different colour, so a wrong row / column / flip / channel order shows).  The
reference reads it with stb_image (image.cpp:11-22, vertical flip on load), the host layer with its own BMP reader."""
import struct
from pathlib import Path

import numpy as np

OUT = Path(__file__).resolve().parent.parent / "assets" / "data" / "tiles.bmp"
nx, ny = 64, 32
y, x = np.mgrid[0:ny, 0:nx]
r = (40 + 3 * x + (y % 4) * 5) % 256
g = (30 + 6 * y + ((x // 8) % 2) * 60) % 256
b = (200 - 2 * x - 3 * y + ((x // 4 + y // 4) % 2) * 50) % 256
rows = np.stack([b, g, r], axis=-1).astype(np.uint8)  # BMP stores BGR, bottom row first
data = rows.tobytes()
hdr = b"BM" + struct.pack("<IHHI", 54 + len(data), 0, 0, 54) + struct.pack("<IiiHHIIiiII", 40, nx, ny, 1, 24, 0, len(data), 2835, 2835, 0, 0)
OUT.write_bytes(hdr + data)
print(OUT, len(hdr) + len(data), "bytes")
