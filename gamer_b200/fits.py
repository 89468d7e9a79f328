"""FITS export of the linear render buffer.

The reference saves one plane per colour channel as BITPIX = -64 (`FitsIO::Savedouble`,
source/util/fitsio.h:18-56, called from mainwindow.cpp:890-892) with rows bottom-up
(`Buffer2D::toQByteArray`, util/buffer2d.cpp:175-185: `Get(j, size-1-i)`). Its byte "flip" swaps groups
of FOUR bytes over the first quarter of an array of 8-byte doubles, so the files it writes are not valid
big-endian float64; its header cards are not padded to 80 columns either. This writer keeps the
reference's content and orientation and writes standard FITS instead of copying those two bugs:
80-column cards, 2880-byte blocks, big-endian IEEE doubles.
"""
import numpy as np

BLOCK = 2880


def _card(key: str, value: str = "", comment: str = "") -> bytes:
    if key == "END":
        s = "END"
    else:
        s = f"{key:<8}= {value:>20}"
        if comment:
            s += f" / {comment}"
    return s[:80].ljust(80).encode("ascii")


def encode_plane(plane: np.ndarray) -> bytes:
    """One 2-D plane (image row 0 at the top, like Buffer2D) -> FITS bytes, rows bottom-up as the reference."""
    plane = np.asarray(plane)
    h, w = plane.shape
    cards = [_card("SIMPLE", "T", "conforms to FITS"), _card("BITPIX", "-64", "IEEE double"), _card("NAXIS", "2"),
             _card("NAXIS1", str(w)), _card("NAXIS2", str(h)), _card("END")]
    head = b"".join(cards)
    head += b" " * (-len(head) % BLOCK)
    data = np.ascontiguousarray(plane[::-1], dtype=">f8").tobytes()
    data += b"\0" * (-len(data) % BLOCK)
    return head + data


def save_double(filename: str, no: int, render_buffer: np.ndarray) -> None:
    """FitsIO::Savedouble(filename, no, buffer): channel `no` of the float RGB render buffer [N,N,3]."""
    with open(filename, "wb") as f:
        f.write(encode_plane(np.asarray(render_buffer)[:, :, no]))


def save_rgb(base: str, render_buffer: np.ndarray):
    """mainwindow.cpp:890-892: <base>_red.fits, _green.fits, _blue.fits."""
    names = [base + s for s in ("_red.fits", "_green.fits", "_blue.fits")]
    for no, name in enumerate(names):
        save_double(name, no, render_buffer)
    return names


def read_plane(filename: str) -> np.ndarray:
    """Minimal reader for the files written above (tests, round trips): returns the plane with row 0 at the top."""
    d = open(filename, "rb").read()
    cards, o = {}, 0
    while True:
        c = d[o:o + 80].decode("ascii")
        o += 80
        if c.startswith("END"):
            break
        k, _, v = c.partition("=")
        cards[k.strip()] = v.split("/")[0].strip()
    o += -o % BLOCK
    assert cards["SIMPLE"] == "T" and cards["BITPIX"] == "-64" and cards["NAXIS"] == "2"
    w, h = int(cards["NAXIS1"]), int(cards["NAXIS2"])
    assert (len(d) - o) % BLOCK == 0 and len(d) - o >= w * h * 8
    return np.frombuffer(d, dtype=">f8", count=w * h, offset=o).reshape(h, w)[::-1].astype(np.float64)
