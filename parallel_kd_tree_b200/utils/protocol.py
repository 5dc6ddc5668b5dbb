"""The reference's stdin/stdout/stderr protocol (Utility.cpp:66-124)."""
from __future__ import annotations

import ctypes as C
import sys

import torch

from .._lib import check, lib, take_string

DEFAULT_DIM = 128          # Utility.cpp:98
DEFAULT_NUM_POINTS = 500000  # Utility.cpp:99
NUM_QUERIES = 10           # kdtree_sequential.cpp:144


def format_float(v: float) -> str:
    """Default ostream formatting of a float (6 significant digits), as the reference prints."""
    buf = C.create_string_buffer(64)
    check(lib().kd_format_float(float(v), buf, 64), "format_float")
    return buf.value.decode()


def result_line(point_id: int, distance: float) -> str:
    """``ID: <id> \\t DISTANCE: <d>`` (Utility.cpp:122-124)."""
    buf = C.create_string_buffer(128)
    check(lib().kd_format_result_line(int(point_id), float(distance), buf, 128), "result_line")
    return buf.value.decode()


def format_point(coords: torch.Tensor, point_id: int) -> str:
    """``Point(ID=…, dimension=…, coordinates=[…])`` including the ``, , ..., `` artefact (Node.cpp:16-28)."""
    c = coords.detach().to("cpu", torch.float32).contiguous()
    out, ln = C.c_void_p(), C.c_size_t()
    check(lib().kd_format_point(c.data_ptr(), c.numel(), int(point_id), C.byref(out), C.byref(ln)), "format_point")
    return take_string(out, ln)


def validate_input(seed: int, dim: int, num_points: int, err=None) -> None:
    """Utility.cpp:66-89: same messages on stderr, exit(1) on invalid values."""
    err = err or sys.stderr
    if seed == 0:
        print("Warning: default value 0 used as seed.", file=err)
    if seed < 0:
        print("Seed has to be larger than 0!", file=err)
        raise SystemExit(1)
    if dim <= 0:
        print("Dimension has to be larger than 0!", file=err)
        raise SystemExit(1)
    if num_points <= 0:
        print("Number of points has to be larger than 0!", file=err)
        raise SystemExit(1)
    print(f"\tUsing seed {seed}", file=err)
    print(f"\tUsing point dimensions {dim}", file=err)
    print(f"\tUsing number of points {num_points}\n", file=err)


def specify_problem_stdin(inp=None, out=None, err=None):
    """Utility.cpp:92-102: READY on stdout, prompt on stderr, seed from stdin, fixed 500000 x 128."""
    inp, out, err = inp or sys.stdin, out or sys.stdout, err or sys.stderr
    print("READY", file=out, flush=True)
    print("Specify seed ", end="", file=err, flush=True)
    tok = ""
    while True:  # operator>> skips whitespace and reads one token
        ch = inp.read(1)
        if ch == "":
            break
        if ch.isspace():
            if tok:
                break
            continue
        tok += ch
    try:
        seed = int(tok)
    except ValueError:
        seed = 0  # a failed extraction leaves 0 in the int
    validate_input(seed, DEFAULT_DIM, DEFAULT_NUM_POINTS, err)
    return seed, DEFAULT_DIM, DEFAULT_NUM_POINTS


def specify_problem_argv(argv, prog="kdtree", out=None, err=None):
    """Utility.cpp:104-120: exactly three arguments SEED DIM_POINTS NUM_POINTS."""
    out, err = out or sys.stdout, err or sys.stderr
    if len(argv) != 3:
        print(f"Usage: {prog} SEED DIM_POINTS  NUM_POINTS", file=err)
        raise SystemExit(1)
    print("READY", file=out, flush=True)
    seed, dim, n = (int(a) for a in argv)
    validate_input(seed, dim, n, err)
    return seed, dim, n
