"""fits.py -- cfitsio-free FITS output for the archive path.

Writes the primary HDU that SnapshotRecorder::write produces through FITSWriter / cfitsio
(reference src/WaterfallBackend.cpp:130-211, src/FITSWriter.cpp:40-168): a FLOAT_IMG
(BITPIX -32) of NAXIS1 = band width, NAXIS2 = rows, with the same keyword set, order, value
types and comments -- CRPIX2 as integer, CRVAL2 as long long milliseconds, CDELT2 as double
(ms per row), CRPIX1 / CRVAL1 / CDELT1 as float -- and big-endian pixels padded to 2880-byte
blocks.  Card formatting follows the FITS fixed format and cfitsio's value conventions (TFLOAT
with 7 significant digits, TDOUBLE with 15, a '.' appended to reals that print without one,
strings blank-padded to 8 characters, values right-justified to column 30).  cfitsio is absent
from this image, so formatting is pinned to the reference's code run over the oracle's cfitsio
shim (tests/golden/*/fits.json), not to a cfitsio binary; `[compress]` output (tile-compressed,
quantised by the library) is unavailable on this box and not produced.
"""
import time as _time
from typing import List, Sequence, Tuple

import numpy as np

BLOCK = 2880
PACKAGE_STRING = "radio-observer 0.5rc1"
PACKAGE_URL = "https://github.com/MLAB-project/radio-observer"


def _pad80(s: str) -> str:
    return s[:80].ljust(80)


def _card_value(key: str, value: str, comment: str = "") -> str:
    c = "%-8.8s= %s" % (key, value.rjust(20))
    if comment:
        c += " / " + comment
    return _pad80(c)


def _card_string(key: str, value: str, comment: str = "") -> str:
    q = "'" + value.replace("'", "''")
    q = q.ljust(9) + "'"
    c = "%-8.8s= %-20s" % (key, q)
    if comment:
        c += " / " + comment
    return _pad80(c)


def _fmt_real(v: float, digits: int) -> str:
    s = "%.*G" % (digits, v)
    if "." not in s and "E" in s:
        s = "%.1E" % v
    if "." not in s and "E" not in s:
        s += "."
    return s


def format_time(sec: int, fmt: str) -> str:
    """WFTime::format for UTC stream times (reference src/WFTime.cpp:25-30: gmtime + strftime)."""
    return _time.strftime(fmt, _time.gmtime(sec))


def snapshot_cards(width: int, length: int, *, origin: str, time_sec: int, time_usec: int,
                   fft_sample_rate: np.float32, left_frequency: np.float32, bin_width: np.float32,
                   git_version: str = "", local_time: str = "", date: str = "") -> List[str]:
    """The header of one snapshot, card for card (without END)."""
    ms = int(np.float64(time_sec) * 1000.0 + np.float64(time_usec) / 1000.0)      # (long long)toMilliseconds()
    cards = [
        _card_value("SIMPLE", "T", "file does conform to FITS standard"),
        _card_value("BITPIX", "-32", "number of bits per data pixel"),
        _card_value("NAXIS", "2", "number of data axes"),
        _card_value("NAXIS1", str(int(width)), "length of data axis 1"),
        _card_value("NAXIS2", str(int(length)), "length of data axis 2"),
        _card_value("EXTEND", "T", "FITS dataset may contain extensions"),
        _pad80("COMMENT   FITS (Flexible Image Transport System) format is defined in 'Astronomy"),
        _pad80("COMMENT   and Astrophysics', volume 376, page 359; bibcode: 2001A&A...376..359H"),
        _pad80("COMMENT File created by %s." % PACKAGE_STRING),
        _pad80("COMMENT Git version: %s" % git_version),
        _pad80("COMMENT See %s." % PACKAGE_URL),
        _card_string("ORIGIN", origin),
    ]
    if date:
        cards.append(_card_string("DATE", date, "file creation date (YYYY-MM-DDThh:mm:ss UT)"))
    cards += [
        _pad80("COMMENT Local time: %s" % local_time),
        _card_string("DATE-OBS", format_time(time_sec, "%Y-%m-%dT%H:%M:%S"), "observation date (UTC)"),
        _card_string("CTYPE2", "TIME", "in seconds"),
        _card_value("CRPIX2", "1"),
        _card_value("CRVAL2", str(ms), "unix time of the first FFT row in this file in ms"),
        _card_value("CDELT2", _fmt_real(float(np.float64(1000.0) / np.float64(np.float32(fft_sample_rate))), 15),
                    "time difference between two FFT samples in ms"),
        _card_string("CTYPE1", "FREQ", "in Hz"),
        _card_value("CRPIX1", _fmt_real(1.0, 7)),
        _card_value("CRVAL1", _fmt_real(float(np.float32(left_frequency)), 7),
                    "frequency, in Hz, of the leftmost pixel in the image"),
        _card_value("CDELT1", _fmt_real(float(np.float32(bin_width)), 7),
                    "frequency difference between two neighbouring pixels in Hz"),
    ]
    return cards


def raw_cards(length: int, *, origin: str, time_sec: int, time_usec: int, sample_rate: int,
              git_version: str = "", local_time: str = "", date: str = "") -> List[str]:
    """The header of a bolid's raw-I/Q file (SnapshotRecorder::writeRaw, reference
    src/WaterfallBackend.cpp:214-267): a 2 x length float image, CDELT2 = (float)1000 / (float)sample rate
    as a float card, CRVAL1 / CDELT1 / CRPIX2 as integers."""
    ms = int(np.float64(time_sec) * 1000.0 + np.float64(time_usec) / 1000.0)
    cards = [
        _card_value("SIMPLE", "T", "file does conform to FITS standard"),
        _card_value("BITPIX", "-32", "number of bits per data pixel"),
        _card_value("NAXIS", "2", "number of data axes"),
        _card_value("NAXIS1", "2", "length of data axis 1"),
        _card_value("NAXIS2", str(int(length)), "length of data axis 2"),
        _card_value("EXTEND", "T", "FITS dataset may contain extensions"),
        _pad80("COMMENT   FITS (Flexible Image Transport System) format is defined in 'Astronomy"),
        _pad80("COMMENT   and Astrophysics', volume 376, page 359; bibcode: 2001A&A...376..359H"),
        _pad80("COMMENT File created by %s." % PACKAGE_STRING),
        _pad80("COMMENT Git version: %s" % git_version),
        _pad80("COMMENT See %s." % PACKAGE_URL),
        _card_string("ORIGIN", origin),
    ]
    if date:
        cards.append(_card_string("DATE", date, "file creation date (YYYY-MM-DDThh:mm:ss UT)"))
    cards += [
        _pad80("COMMENT Local time: %s" % local_time),
        _card_string("DATE-OBS", format_time(time_sec, "%Y-%m-%dT%H:%M:%S"), "observation date (UTC)"),
        _card_string("CTYPE2", "TIME", "in seconds"),
        _card_value("CRPIX2", "1"),
        _card_value("CRVAL2", str(ms), "unix time of the first IQ sample in this file in ms"),
        _card_value("CDELT2", _fmt_real(float(np.float32(np.float32(1000.0) / np.float32(sample_rate))), 7),
                    "time difference between two IQ samples in ms"),
        _card_string("CTYPE1", "CHAN", "in Hz"),
        _card_value("CRPIX1", _fmt_real(1.0, 7)),
        _card_value("CRVAL1", "0"),
        _card_value("CDELT1", "1"),
    ]
    return cards


def file_basename(time_sec: int, time_usec: int, origin: str, typ: str, ext: str = "fits") -> str:
    """SnapshotRecorder::getFileBasename (reference src/WaterfallBackend.cpp:320-336)."""
    return "%s%03d_%s_%s.%s" % (format_time(time_sec, "%Y%m%d%H%M%S"), time_usec // 1000, origin, typ, ext)


def write_image(path: str, cards: Sequence[str], data: np.ndarray) -> None:
    """Header cards + END, then `data` ([rows][width] float32) big-endian, each padded to 2880 bytes."""
    hdr = "".join(_pad80(c) for c in cards) + _pad80("END")
    hdr = hdr.ljust(-(-len(hdr) // BLOCK) * BLOCK)
    px = np.ascontiguousarray(data, dtype=">f4").tobytes()
    px += b"\0" * (-len(px) % BLOCK)
    with open(path, "wb") as f:
        f.write(hdr.encode("ascii"))
        f.write(px)


def read_image(path: str) -> Tuple[List[str], dict, np.ndarray]:
    """(cards, {keyword: value text}, data[rows][width]) of a primary-HDU float image."""
    b = open(path, "rb").read()
    cards, pos, end = [], 0, False
    while not end:
        block = b[pos:pos + BLOCK]
        if len(block) < BLOCK:
            raise ValueError("%s: truncated FITS header" % path)
        pos += BLOCK
        for i in range(0, BLOCK, 80):
            c = block[i:i + 80].decode("ascii")
            if c.startswith("END"):
                end = True
                break
            cards.append(c.rstrip())
    hdr = {}
    for c in cards:
        if c[8:10] == "= ":
            hdr[c[:8].strip()] = c[10:].split(" /")[0].strip()
    if int(hdr.get("BITPIX", "0")) != -32 or int(hdr.get("NAXIS", "0")) != 2:
        raise ValueError("%s: not a 2-D float image" % path)
    w, h = int(hdr["NAXIS1"]), int(hdr["NAXIS2"])
    data = np.frombuffer(b, dtype=">f4", count=w * h, offset=pos).reshape(h, w)
    return cards, hdr, data
