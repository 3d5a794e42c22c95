"""ORACLE — test infrastructure only.  Never imported by the nums_b200 package.

A restatement of the reference's CSV block reader (nums/core/filesystem.py:227-283, v0.2.8) on the bytes of a file:
which lines a (file_start, file_end) byte range owns and how each field is converted.  Pinned by
tests/test_oracle_golden.py against tests/golden/csv.npz, produced by calling the reference's own read_csv_block
(tests/golden/make_golden.py).
"""
import numpy as np


def _converter(dtype):
    """filesystem.py:228-254 (`_getconv`) for the dtypes the ingest path takes."""
    def floatconv(x):
        if "0x" in x:
            return float.fromhex(x)
        return float(x)
    if dtype is int:
        return lambda x: int(float(x))
    if issubclass(dtype, np.int64):
        return np.int64
    return floatconv


def read_csv_block(data, file_start, file_end, dtype=float, delimiter=",", has_header=False):
    """data: the whole file as bytes (ASCII).  Returns the block's array (shape (0,) when it owns no line)."""
    # text mode with universal newlines: "\r\n" reads as "\n" (the "\r" is stripped below either way), offsets are
    # byte offsets for ASCII content; a lone "\r" as a line terminator is not restated
    text = data.decode("utf-8")
    pos = file_start
    if file_start != 0:
        k = text.find("\n", pos)
        pos = len(text) if k < 0 else k + 1
    conv = _converter(dtype)
    lines, header_skipped = [], False
    while pos < file_end and pos < len(text):
        k = text.find("\n", pos)
        line = text[pos:] if k < 0 else text[pos:k + 1]
        pos = len(text) if k < 0 else k + 1
        line = line.strip("\r\n")
        if file_start == 0 and has_header and not header_skipped:
            header_skipped = True
            continue
        lines.append([conv(f) for f in line.split(delimiter)])
    return np.array(lines, dtype=dtype)


def batches(total_size, num_batches):
    """storage/utils.py:29-61"""
    batch_size = (total_size + num_batches - 1) // num_batches
    if total_size < batch_size:
        return [[0, total_size]]
    starts = list(range(0, total_size, batch_size))
    out = [starts[i:i + 2] for i in range(int(total_size / batch_size))]
    if len(out[-1]) == 1:
        out[-1].append(total_size)
    if out[-1][1] != total_size:
        out.append([out[-1][1], total_size])
    return out


def read_csv(data, dtype=float, delimiter=",", has_header=False, num_workers=4):
    """filesystem.py:615-657 + application.py:172-190: the concatenation of the non-empty blocks."""
    parts = [read_csv_block(data, s, e, dtype, delimiter, has_header) for s, e in batches(len(data), num_workers)]
    parts = [p for p in parts if p.shape[0] > 0]
    return np.concatenate(parts, axis=0)
