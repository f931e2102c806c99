#!/usr/bin/env python3
"""tsbin -- codec between the ASCII / gzip series files and the raw binary files the host driver reads and writes.

Mirrors the reference tool util/tsbin.py (command line :113-190, formats :18-112) so that the datasets shipped with
the reference (data/real/*.txt, data/synthetic/*.txt, data/benchmark/*.txt.gz, data/README.md:5-14) run through
MatrixProfileHost and its results can be read back:

* binary  : headerless little-endian array, ``double`` = float64 (.tsb series, .mpb profile) or ``int`` = int32
            (.mpib index) -- include/host/FileIO.hpp:24-48,105-121;
* ASCII   : whitespace-separated numbers, one per line or all on one line, optionally gzip-compressed (.txt.gz).

Same flags and rules as the reference: -e/-d (exactly one), -o (mandatory for -e), -t {double,int} (mandatory for -e,
default double for -d), -n (expected element count), -l/--limit, --offset, --oneline (decode only), -v.  Errors are
reported as ``[ERROR] ...`` and a non-zero exit status.  Arrays go through numpy instead of struct.pack of Python lists,
so the n = 2^24 series of the large configurations encode in a second instead of minutes.
"""
import argparse
import gzip
import os
import sys

import numpy as np

TYPES = {"double": (np.dtype("<f8"), float), "int": (np.dtype("<i4"), int)}
DEFAULT_TYPE = "double"
_verbose = False


class TsbinError(Exception):
    pass


def log(msg, level="INFO"):
    if level != "INFO/V" or _verbose:
        print(f"[{level}] {msg}", file=sys.stderr if level == "ERROR" else sys.stdout)


def _dtype(type_):
    if type_ is None:
        log(f"No type specified defaulting to {DEFAULT_TYPE}", "INFO/V")
        type_ = DEFAULT_TYPE
    if type_ not in TYPES:
        raise TsbinError(f'Unknown type "{type_}". Type has to be one of: {", ".join(TYPES)}')
    return type_, TYPES[type_][0]


def read_bin(filename, type_=None, n=None):
    """Binary file -> 1-D numpy array (util/tsbin.py:18-52): the size must be a multiple of the element size and,
    when n is given, exactly n elements."""
    type_, dt = _dtype(type_)
    try:
        size = os.path.getsize(filename)
    except OSError:
        raise TsbinError(f'Cannot determine size of file "{filename}"')
    if n is None:
        if size % dt.itemsize != 0:
            raise TsbinError(f"File contains {size} bytes, which is not a multiple of {dt.itemsize} bytes (type = {type_})")
        n = size // dt.itemsize
        log(f"Determined number of elements (n = {n})", "INFO/V")
    if size != dt.itemsize * n:
        raise TsbinError(f"{filename} contains unexpected number of elements! Expected {n} element(s) [i.e. {n * dt.itemsize} bytes]; but file contains {size} bytes")
    try:
        data = np.fromfile(filename, dtype=dt, count=n)
    except OSError:
        raise TsbinError(f'Cannot read file "{filename}"')
    log(f"Read and decoded file \"{filename}\" ({size} bytes, i.e. {n} elements of type '{type_}')", "INFO/V")
    return data


def write_bin(filename, data, type_):
    """1-D array -> headerless little-endian binary (util/tsbin.py:54-68)."""
    type_, dt = _dtype(type_)
    arr = np.ascontiguousarray(np.asarray(data), dtype=dt)
    try:
        arr.tofile(filename)
    except OSError:
        raise TsbinError(f'Cannot write file "{filename}"')
    log(f"Wrote and encoded file \"{filename}\" ({arr.nbytes} bytes, i.e. {arr.size} elements of type '{type_}')")
    return True


def read_ascii(filename, type_=None):
    """.txt / .txt.gz -> 1-D array (util/tsbin.py:70-99): numbers split on any whitespace, blank lines ignored, one
    number per line and all-on-one-line both accepted; ``int`` parses with int() semantics (a decimal point is an error)."""
    type_, dt = _dtype(type_)
    try:
        if filename.endswith(".txt.gz"):
            with gzip.open(filename, "rb") as f:
                text = f.read().decode("utf-8")
        elif filename.endswith(".txt"):
            with open(filename, "r") as f:
                text = f.read()
        else:
            raise OSError("unsupported extension")
    except (OSError, UnicodeDecodeError, EOFError):
        raise TsbinError(f'Cannot read file "{filename}"')
    tokens = text.split()
    try:
        if type_ == "int":
            data = np.array([int(t) for t in tokens], dtype=np.int64)
            if data.size and (data.min() < -2**31 or data.max() >= 2**31):
                raise ValueError("out of int32 range")
            data = data.astype(dt)
        else:
            data = np.array(tokens, dtype=np.float64) if tokens else np.empty(0, np.float64)   # same correctly-rounded parse as float()
    except ValueError:
        raise TsbinError(f'Cannot decode file "{filename}"')
    return data


def _fmt(x, type_):
    return repr(float(x)) if type_ == "double" else str(int(x))      # str(float) == repr(float): shortest round-trip form


def write_ascii(filename, data, type_=None, oneline=False):
    """1-D array -> text, one number per line or space-separated on one line (util/tsbin.py:101-111)."""
    type_, _ = _dtype(type_)
    try:
        with open(filename, "w") as f:
            f.write((" " if oneline else "\n").join(_fmt(x, type_) for x in data) + "\n")
    except OSError:
        raise TsbinError(f'Cannot write file "{filename}"')
    log(f"Wrote file \"{filename}\" ({len(data)} elements of type '{type_}')")
    return True


def _window(data, offset, limit):
    """--offset / --limit as the reference applies them: data[offset : offset + limit] (:158-168)."""
    offset = offset or 0
    end = offset + limit if limit else len(data)
    return data[offset:end]


def main(argv=None):
    global _verbose
    p = argparse.ArgumentParser(description="encode/decode binary time series files", usage="tsbin.py [OPTIONS]")
    g = p.add_mutually_exclusive_group()
    g.add_argument("-d", "--decode", action="store_true", help="decode file from binary")
    g.add_argument("-e", "--encode", action="store_true", help="encode file to binary")
    p.add_argument("input", nargs=1, help="the file to encode/decode")
    p.add_argument("-o", "--output", help="specify the output file name")
    p.add_argument("-t", "--type", choices=list(TYPES), help="type of numbers")
    p.add_argument("-n", type=int, help="count of numbers to be decode [if -n is not specified value will be inferred]")
    p.add_argument("-l", "--limit", type=int, help="limit the count of numbers to encode/decode")
    p.add_argument("--offset", type=int, help="ignore the first OFFSET numbers when encoding/decoding")
    p.add_argument("--oneline", action="store_true", help="output decoded file in a single line")
    p.add_argument("-v", "--verbose", action="store_true", help="increase output verbosity")
    a = p.parse_args(argv)
    _verbose = bool(a.verbose)
    for name, v in (("-n", a.n), ("-l/--limit", a.limit), ("--offset", a.offset)):
        if v is not None and v < 0:
            p.error(f"{name} must have a non-negative value")
    try:
        if a.encode:
            if not a.output:
                p.error("-o/--output has to be specified in -e/--encode mode")
            if not a.type:
                p.error("-t/--type has to be specified in -e/--encode mode")
            if a.oneline:
                p.error("--oneline is not applicate in -e/--encode mode")
            data = read_ascii(a.input[0], a.type)
            if a.n and len(data) != a.n:
                raise TsbinError(f"Failed to decode {a.n} numbers. Decoded {len(data)} instead [aborting]")
            write_bin(a.output, _window(data, a.offset, a.limit), a.type)
        elif a.decode:
            data = _window(read_bin(a.input[0], a.type, a.n), a.offset, a.limit)
            type_ = a.type or DEFAULT_TYPE
            if not a.output:
                print(*(_fmt(x, type_) for x in data), sep=(", " if a.oneline else "\n"))
            else:
                write_ascii(a.output, data, type_, a.oneline)
        else:
            p.error("either -d/--decode or -e/--encode has to be set")
    except TsbinError as e:
        log(str(e), "ERROR")
        return 1
    return 0


if __name__ == "__main__":
    sys.exit(main())
