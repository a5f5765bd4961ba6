"""MPS reader: host-side restatement of the reference's `io::import`
(/root/reference/src/io/mod.rs:32-64) -- free format (`parse/free.rs`, whitespace tokens) and fixed
format (`parse/fixed.rs:120-128`, column ranges) over one section parser (`parse/mod.rs:37-90`),
exact decimal numbers (`number/parse.rs:44-121`: every number becomes an exact rational, no floats).

Semantics kept from the reference:
  * rows are SORTED BY NAME after the ROWS section (`parse/mod.rs` check_row_section_consistency);
    columns keep file order; a column's entries are sorted by row index;
  * comment lines start with `*`; empty lines are skipped; OBJSENSE (next line MAX/MIN) is optional;
  * MARKER lines switch integrality (ignored by the LP relaxation);
  * RHS / RANGES / BOUNDS groups are merged regardless of their names.
Deviation (documented): an RHS entry on the objective row is an error in the reference ("Row not
known"); here it is accepted as minus the objective constant when `strict=False`.
"""
import re
from fractions import Fraction as Fr

FIXED_FIELDS = [(0, 1), (1, 3), (4, 12), (14, 22), (24, 36), (39, 47), (49, 61)]   # parse/fixed.rs:120-128
SECTIONS = ("NAME", "OBJSENSE", "ROWS", "COLUMNS", "RHS", "RANGES", "BOUNDS", "ENDATA")


class MPSError(ValueError):
    pass


class MPS:
    """`io::mps::MPS<Rational64>` (mps/mod.rs): raw sections, indices already resolved."""

    def __init__(self):
        self.name = ""
        self.maximize = False
        self.cost_row_name = None
        self.rows = []            # (name, 'E'|'L'|'G'), sorted by name
        self.columns = []         # (name, [(row, Fr)] sorted by row, is_integer)
        self.cost = {}            # column index -> Fr
        self.rhs = []             # [(row, Fr)] all groups, file order
        self.ranges = []          # [(row, Fr)]
        self.bounds = []          # [(column, type, Fr | None)] file order
        self.objective_constant = Fr(0)


_I64_MAX = 2 ** 63 - 1
_RUST_INT = re.compile(r"[+-]?[0-9]+\Z")


def _rust_i64(text, part):
    """`text.parse::<i64>()` as number/parse.rs:98-108 uses it: an empty part is 0; otherwise ASCII digits with one
    optional sign, within the i64 range -- nothing else (no exponent, no '/', no '_', no blanks)."""
    if text == "":
        return 0
    if not _RUST_INT.match(text):
        raise MPSError(f"Failed to parse {part} {text!r} as i64")
    v = int(text)
    if not -_I64_MAX - 1 <= v <= _I64_MAX:
        raise MPSError(f"{part} {text!r} does not fit i64")
    return v


def parse_raw(text):
    """`Raw::try_from` (number/parse.rs:88-121): (negative?, integer, decimal_steps_from_right).  Refuses exactly
    what the reference refuses (scientific notation, fractions, separators, numbers beyond i64)."""
    if text == "":
        raise MPSError("empty number")                      # the reference slices text[0..1] and panics
    negative = text[0] == "-"
    if negative:
        text = text[1:]
    dot = text.find(".")
    if dot < 0:
        return negative, _rust_i64(text, "entire value"), 0
    from_right = len(text) - dot - 1
    if from_right > 18:
        raise MPSError(f"{text!r}: 10^{from_right} does not fit i64")
    integer = _rust_i64(text[:dot], "integer part") * 10 ** from_right + _rust_i64(text[dot + 1:], "mantissa part")
    if not -_I64_MAX - 1 <= integer <= _I64_MAX:
        raise MPSError(f"{text!r} does not fit i64")
    return negative, integer, from_right


def parse_number(text):
    """`<Rational64 as Parse>::parse` (number/parse.rs:44-67): exact decimal -> reduced rational."""
    negative, integer, from_right = parse_raw(text.strip())
    return Fr(-integer if negative else integer, 10 ** from_right)


def _fields(line, fixed):
    if not fixed:
        return line.split()
    out = []
    for k, (a, b) in enumerate(FIXED_FIELDS):
        if k == 0:
            continue
        s = line[a:b].strip()
        if s:
            out.append(s)
    return out


def parse(text, fixed=False, strict=False):
    """`mps::parse` (free, io/mps/mod.rs:36) / `mps::parse_fixed` (:56)."""
    mps = MPS()
    section = None
    rows_raw, row_index = [], {}
    col_index = {}
    integer = False
    cur_col, cur_vals = None, None
    expect_objsense = False

    def finish_column():
        nonlocal cur_col, cur_vals
        if cur_col is not None:
            cur_vals.sort(key=lambda x: x[0])
            for a, b in zip(cur_vals, cur_vals[1:]):
                if a[0] == b[0]:
                    raise MPSError(f"Duplicate row for column {cur_col!r}")
            col_index[cur_col] = len(mps.columns)
            mps.columns.append((cur_col, cur_vals, integer_of_col[0]))
        cur_col, cur_vals = None, None

    integer_of_col = [False]
    for lineno, raw in enumerate(text.splitlines(), 1):
        line = raw.rstrip("\r\n")
        if not line.strip() or line.lstrip().startswith("*"):
            continue
        if not line[0].isspace():                      # section header
            head = line.split()
            key = head[0].upper()
            if key not in SECTIONS:
                raise MPSError(f"line {lineno}: unknown section {head[0]!r}")
            if section == "ROWS" and key != "ROWS":
                # rows sorted by name; build the index (parse/mod.rs check_row_section_consistency)
                if mps.cost_row_name is None:
                    raise MPSError("No cost name read.")
                rows_raw.sort(key=lambda r: r[0])
                for a, b in zip(rows_raw, rows_raw[1:]):
                    if a[0] == b[0]:
                        raise MPSError(f"Duplicate row name {a[0]}")
                mps.rows = rows_raw
                row_index = {name: i for i, (name, _) in enumerate(rows_raw)}
            if section == "COLUMNS" and key != "COLUMNS":
                finish_column()
            section = key
            if key == "NAME":
                mps.name = head[1] if len(head) > 1 else ""
            elif key == "OBJSENSE":
                if len(head) > 1:
                    mps.maximize = head[1].upper().startswith("MAX")
                else:
                    expect_objsense = True
            elif key == "ENDATA":
                break
            continue
        if section == "OBJSENSE" and expect_objsense:
            mps.maximize = line.split()[0].upper().startswith("MAX")
            expect_objsense = False
            continue
        f = _fields(line, fixed)
        if section == "ROWS":
            kind, name = f[0].upper(), f[1]
            if kind == "N":
                if mps.cost_row_name is not None:
                    if strict:
                        raise MPSError("Second cost row detected. This is not supported.")
                    continue                                  # further free rows are dropped
                mps.cost_row_name = name
            elif kind in ("E", "L", "G"):
                rows_raw.append((name, kind))
            else:
                raise MPSError(f"line {lineno}: row type {kind!r}")
        elif section == "COLUMNS":
            if len(f) >= 3 and f[1].strip("'").upper() == "MARKER":
                finish_column()
                integer = "INTORG" in f[2].upper()
                continue
            name = f[0]
            if name != cur_col:
                finish_column()
                if name in col_index:
                    raise MPSError(f"column {name!r} is not contiguous")
                cur_col, cur_vals = name, []
                integer_of_col = [integer]
            for k in range(1, len(f) - 1, 2):
                rname, val = f[k], parse_number(f[k + 1])
                if rname == mps.cost_row_name:
                    mps.cost[len(mps.columns)] = val
                elif rname in row_index:
                    cur_vals.append((row_index[rname], val))
                else:
                    raise MPSError(f'Row "{rname}" not known.')
        elif section in ("RHS", "RANGES"):
            # group name may be missing in free format when the count of fields is even
            start = 1 if len(f) % 2 == 1 else 0
            for k in range(start, len(f) - 1, 2):
                rname, val = f[k], parse_number(f[k + 1])
                if rname in row_index:
                    (mps.rhs if section == "RHS" else mps.ranges).append((row_index[rname], val))
                elif rname == mps.cost_row_name and section == "RHS" and not strict:
                    mps.objective_constant = -val
                else:
                    raise MPSError(f'Row "{rname}" not known.')
        elif section == "BOUNDS":
            kind = f[0].upper()
            if kind in ("FR", "MI", "PL", "BV"):
                cname = f[2] if len(f) >= 3 else f[1]
                val = None
            else:
                if len(f) >= 4:
                    cname, val = f[2], parse_number(f[3])
                else:                                         # bound name omitted
                    cname, val = f[1], parse_number(f[2])
            if cname not in col_index:
                raise MPSError(f'Column name "{cname}" unknown')
            if kind not in ("FR", "MI", "PL", "BV", "LO", "UP", "FX", "LI", "UI"):
                raise MPSError(f'Bound type "{kind}" unknown.')
            mps.bounds.append((col_index[cname], kind, val))
        else:
            raise MPSError(f"line {lineno}: data before any section")
    if section == "COLUMNS":
        finish_column()
    dup = set()
    for i, _ in mps.ranges:
        if i in dup:
            raise MPSError("Each row can have at most one range value")
        dup.add(i)
    return mps


def import_file(path, fixed=None, strict=False):
    """`io::import` (io/mod.rs:32-64): `.mps` -> free format, `.SIF` -> the netlib tests use the fixed
    parser (tests/netlib/mod.rs:54)."""
    if fixed is None:
        fixed = path.lower().endswith(".sif")
    with open(path) as fh:
        return parse(fh.read(), fixed=fixed, strict=strict)
