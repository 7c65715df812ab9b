"""Minimal FITS binary-table reader / writer (numpy only).

Init-time I/O for the hot path's static inputs: the reference reads its
correlation files, distortion matrices, covariance, metal matrices and P(k)
templates through ``astropy.io.fits`` (vega/data.py:302-394,450-473,578-628;
vega/vega_interface.py:690-701,899-900).  astropy is not in this image, so this
module implements the small subset those call sites touch: BINTABLE HDUs with
scalar and fixed-width vector columns (TFORM ``D E K J I B L A`` with repeat
counts), header cards, gzip transparently, lookup by index or EXTNAME.

Nothing here runs per chi2 evaluation.
"""
import builtins
import gzip

import numpy as np

_BLOCK = 2880

_TFORM_DTYPES = {
    'D': '>f8', 'E': '>f4', 'K': '>i8', 'J': '>i4', 'I': '>i2',
    'B': 'u1', 'L': 'S1', 'A': 'S',
}


def _parse_value(raw):
    """Parse the value field of a header card."""
    raw = raw.strip()
    if not raw:
        return None
    if raw.startswith("'"):
        end = 1
        out = []
        while end < len(raw):
            ch = raw[end]
            if ch == "'":
                if end + 1 < len(raw) and raw[end + 1] == "'":
                    out.append("'")
                    end += 2
                    continue
                break
            out.append(ch)
            end += 1
        return ''.join(out).rstrip()
    raw = raw.split('/')[0].strip()
    if raw == 'T':
        return True
    if raw == 'F':
        return False
    try:
        return int(raw)
    except ValueError:
        pass
    try:
        return float(raw.replace('D', 'E'))
    except ValueError:
        return raw


class Header(dict):
    """Header cards as a dict (keeps insertion order)."""

    def get(self, key, default=None):
        return dict.get(self, key, default)


class _Columns:
    def __init__(self, names):
        self.names = list(names)


class _TableData:
    """Column access by name returning native-endian arrays."""

    def __init__(self, columns):
        self._columns = columns

    def __getitem__(self, name):
        return self._columns[name]

    def __contains__(self, name):
        return name in self._columns

    def keys(self):
        return self._columns.keys()


class HDU:
    def __init__(self, header, columns=None):
        self.header = header
        self.name = header.get('EXTNAME', '')
        if columns is None:
            self.data = None
            self.columns = _Columns([])
        else:
            self.data = _TableData(columns)
            self.columns = _Columns(columns.keys())


class HDUList(list):
    def __getitem__(self, key):
        if isinstance(key, str):
            for hdu in self:
                if hdu.name == key:
                    return hdu
            raise KeyError(f'Extension {key} not found')
        return list.__getitem__(self, key)

    def close(self):
        pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


def _read_header(buf, pos):
    header = Header()
    while True:
        block = buf[pos:pos + _BLOCK]
        if len(block) < _BLOCK:
            raise EOFError
        pos += _BLOCK
        finished = False
        for i in range(36):
            card = block[i * 80:(i + 1) * 80].decode('ascii', errors='replace')
            key = card[:8].strip()
            if key == 'END':
                finished = True
                break
            if key == 'HIERARCH' and '=' in card:
                name, _, value = card[9:].partition('=')
                header[name.strip()] = _parse_value(value)
                continue
            if key in ('', 'COMMENT', 'HISTORY') or card[8:10] != '= ':
                continue
            header[key] = _parse_value(card[10:])
        if finished:
            return header, pos


def _split_tform(tform):
    tform = tform.strip()
    i = 0
    while i < len(tform) and tform[i].isdigit():
        i += 1
    repeat = int(tform[:i]) if i > 0 else 1
    return repeat, tform[i]


def open(path, *args, **kwargs):  # noqa: A001 - mirrors astropy.io.fits.open
    """Read every HDU of a (possibly gzipped) FITS file into memory."""
    path = str(path)
    with builtins.open(path, 'rb') as f:
        magic = f.read(2)
    opener = gzip.open if magic == b'\x1f\x8b' else builtins.open
    with opener(path, 'rb') as f:
        buf = f.read()

    hdus = HDUList()
    pos = 0
    while pos < len(buf):
        try:
            header, pos = _read_header(buf, pos)
        except EOFError:
            break
        naxis = header.get('NAXIS', 0)
        nbytes = 0
        if naxis:
            nbytes = abs(header['BITPIX']) // 8
            for ax in range(1, naxis + 1):
                nbytes *= header[f'NAXIS{ax}']
            nbytes += header.get('PCOUNT', 0)

        if header.get('XTENSION', '') == 'BINTABLE':
            nrows = header['NAXIS2']
            fields = []
            names = []
            for i in range(1, header['TFIELDS'] + 1):
                repeat, code = _split_tform(header[f'TFORM{i}'])
                name = header.get(f'TTYPE{i}', f'col{i}')
                names.append(name)
                if code == 'A':
                    fields.append((name, f'S{repeat}'))
                elif repeat == 1:
                    fields.append((name, _TFORM_DTYPES[code]))
                else:
                    fields.append((name, _TFORM_DTYPES[code], (repeat,)))
            dtype = np.dtype(fields)
            assert dtype.itemsize == header['NAXIS1'], 'Unsupported table layout'
            table = np.frombuffer(buf, dtype=dtype, count=nrows, offset=pos)
            columns = {}
            for name in names:
                col = table[name]
                if col.dtype.kind in 'fiu':
                    col = col.astype(col.dtype.newbyteorder('='))
                elif col.dtype.kind == 'S' and col.dtype.itemsize == 1 \
                        and header[f'TFORM{names.index(name) + 1}'].strip().endswith('L'):
                    col = (col == b'T')
                else:
                    col = np.char.strip(col.astype(str))
                columns[name] = col
            hdus.append(HDU(header, columns))
        else:
            hdus.append(HDU(header))

        pos += ((nbytes + _BLOCK - 1) // _BLOCK) * _BLOCK
    return hdus


# ----------------------------------------------------------------------------
# Writer (used to produce synthetic DESI/eBOSS-shaped inputs and fixtures)
# ----------------------------------------------------------------------------
def _card(key, value, comment=''):
    key = str(key)
    if key.lower().startswith('hierarch '):
        key = key[9:]
    if len(key) > 8 or not all(ch.isupper() or ch.isdigit() or ch in '-_' for ch in key):
        # ESO HIERARCH convention for long / mixed-case keywords (what astropy writes for them)
        if isinstance(value, (bool, np.bool_)):
            val = 'T' if value else 'F'
        elif isinstance(value, (int, np.integer)):
            val = str(int(value))
        elif isinstance(value, (float, np.floating)):
            val = repr(float(value))
        else:
            val = "'" + str(value).replace("'", "''") + "'"
        return f'HIERARCH {key} = {val}'[:80].ljust(80)
    if isinstance(value, (bool, np.bool_)):
        val = 'T' if value else 'F'
        text = f'{key:<8}= {val:>20}'
    elif isinstance(value, (int, np.integer)):
        text = f'{key:<8}= {int(value):>20d}'
    elif isinstance(value, (float, np.floating)):
        text = f'{key:<8}= {repr(float(value)):>20}'
    else:
        sval = "'" + str(value).replace("'", "''").ljust(8) + "'"
        text = f'{key:<8}= {sval:<20}'
    if comment:
        text += f' / {comment}'
    return text[:80].ljust(80)


def _header_bytes(cards):
    text = ''.join(cards) + 'END'.ljust(80)
    pad = (-len(text)) % _BLOCK
    return (text + ' ' * pad).encode('ascii')


def _table_bytes(columns, extname, extra_header):
    names = list(columns.keys())
    fields = []
    tforms = []
    nrows = None
    for name in names:
        arr = np.asarray(columns[name])
        if nrows is None:
            nrows = arr.shape[0]
        assert arr.shape[0] == nrows, 'All columns need the same number of rows'
        repeat = int(np.prod(arr.shape[1:])) if arr.ndim > 1 else 1
        if arr.dtype.kind == 'f':
            code, dt = 'D', '>f8'
        elif arr.dtype.kind in 'iu':
            code, dt = 'K', '>i8'
        elif arr.dtype.kind == 'b':
            code, dt = 'L', 'S1'
        elif arr.dtype.kind in 'US':
            width = max(1, max((len(str(x)) for x in arr.ravel()), default=1))
            tforms.append(f'{width}A')
            fields.append((name, f'S{width}'))
            continue
        else:
            raise TypeError(f'Unsupported column dtype {arr.dtype}')
        tforms.append(f'{repeat}{code}' if repeat > 1 else code)
        fields.append((name, dt) if repeat == 1 else (name, dt, (repeat,)))
    dtype = np.dtype(fields)
    table = np.zeros(nrows, dtype=dtype)
    for name in names:
        arr = np.asarray(columns[name])
        if arr.dtype.kind == 'b':
            arr = np.where(arr, b'T', b'F')
        elif arr.dtype.kind == 'U':
            arr = np.char.encode(arr, 'ascii')
        table[name] = arr.reshape(nrows, -1) if arr.ndim > 1 else arr

    cards = [
        _card('XTENSION', 'BINTABLE'), _card('BITPIX', 8), _card('NAXIS', 2),
        _card('NAXIS1', dtype.itemsize), _card('NAXIS2', nrows), _card('PCOUNT', 0),
        _card('GCOUNT', 1), _card('TFIELDS', len(names)),
    ]
    for i, (name, tform) in enumerate(zip(names, tforms), start=1):
        cards.append(_card(f'TTYPE{i}', name))
        cards.append(_card(f'TFORM{i}', tform))
    if extname:
        cards.append(_card('EXTNAME', extname))
    for key, value in (extra_header or {}).items():
        cards.append(_card(key, value))

    raw = table.tobytes()
    pad = (-len(raw)) % _BLOCK
    return _header_bytes(cards) + raw + b'\0' * pad


def write_tables(path, tables, compress=None):
    """Write a FITS file made of an empty primary HDU and binary tables.

    Parameters
    ----------
    path : str
        Output path; ``.gz`` suffix (or ``compress=True``) gzips the file
    tables : list of (extname, header_dict, columns_dict)
        Each column is a 1-D array (scalar column) or 2-D array (vector column)
    """
    path = str(path)
    primary = _header_bytes([
        _card('SIMPLE', True), _card('BITPIX', 8), _card('NAXIS', 0), _card('EXTEND', True)])
    payload = primary + b''.join(
        _table_bytes(columns, extname, header) for extname, header, columns in tables)
    if compress is None:
        compress = path.endswith('.gz')
    if compress:
        with gzip.open(path, 'wb', compresslevel=6) as f:
            f.write(payload)
    else:
        with builtins.open(path, 'wb') as f:
            f.write(payload)
