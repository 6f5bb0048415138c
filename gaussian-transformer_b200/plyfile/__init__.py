"""Minimal `plyfile` stand-in ("+ a minimal plyfile shim").

The reference imports the third-party `plyfile` package (not vendored, not installed in this image) at
`scene/gaussian_model.py:18` and `scene/dataset_readers.py:22`; without it `scene.gaussian_model` cannot be imported and
`gaussian_renderer.render()` / `train.py` cannot run unchanged.  This module covers exactly the calls those two files
make, with plyfile's names, argument meaning and error types:

* `PlyData.read(path)`                       `scene/gaussian_model.py:216`, `scene/dataset_readers.py:108`
* `plydata.elements[0]["x"]`, `plydata['vertex']`, `vertices['red']`   `gaussian_model.py:218-247`, `dataset_readers.py:109-112`
* `plydata.elements[0].properties` → objects with `.name`              `gaussian_model.py:228,237,243`
* `PlyElement.describe(structured_array, 'vertex')`                    `gaussian_model.py:207`, `dataset_readers.py:128`
* `PlyData([el]).write(path)`                                          `gaussian_model.py:208`, `dataset_readers.py:129-130`

Files are the PLY 1.0 format: an ASCII header (`ply`, `format`, `comment`, `element <name> <count>`, `property <type>
<name>`, `end_header`) followed by the elements in order, as text or as packed little/big-endian records.  Written files
use plyfile's defaults (binary, native byte order, type names `float`/`uchar`/...), so that a file saved through this
shim is byte-identical to the one the real package writes for the same array.  List properties (`property list uchar int
vertex_indices`, the faces of a mesh) do not occur on this path -- both call sites read and write one `vertex` element of
scalars -- but are read and written too (object columns of numpy arrays, row by row), so that a mesh PLY handed to
`fetchPly` loads its vertices instead of failing.
"""
import sys

import numpy as np

__all__ = ["PlyData", "PlyElement", "PlyProperty", "PlyListProperty", "PlyParseError", "PlyHeaderParseError", "PlyElementParseError"]

# PLY type name -> numpy kind+size.  Both the classic names and the sized aliases are legal in headers.
_TYPES = {
    "char": "i1", "uchar": "u1", "short": "i2", "ushort": "u2", "int": "i4", "uint": "u4", "float": "f4", "double": "f8",
    "int8": "i1", "uint8": "u1", "int16": "i2", "uint16": "u2", "int32": "i4", "uint32": "u4", "float32": "f4",
    "float64": "f8",
}
# numpy kind+size -> the name written into headers (the classic one, as plyfile does).
_NAMES = {"i1": "char", "u1": "uchar", "i2": "short", "u2": "ushort", "i4": "int", "u4": "uint", "f4": "float",
          "f8": "double"}
_FORMATS = {"ascii": "=", "binary_little_endian": "<", "binary_big_endian": ">"}
_NATIVE = "<" if sys.byteorder == "little" else ">"


class PlyParseError(Exception):
    """Base class of the parse errors, as in plyfile."""


class PlyHeaderParseError(PlyParseError):
    def __init__(self, message, line=None):
        self.message, self.line = message, line
        super().__init__(message if line is None else "line %d: %s" % (line, message))


class PlyElementParseError(PlyParseError):
    def __init__(self, message, element=None, row=None, prop=None):
        self.message, self.element, self.row, self.prop = message, element, row, prop
        where = ""
        if element is not None:
            where += "element %r: " % element.name
        if row is not None:
            where += "row %d: " % row
        super().__init__(where + message)


class PlyProperty:
    """One scalar property of an element: `.name`, `.val_dtype` (numpy code without byte order), `.dtype(byte_order)`."""

    def __init__(self, name, val_dtype):
        name = str(name)
        if not name or any(c.isspace() for c in name):
            raise ValueError("invalid property name %r" % name)
        code = _TYPES.get(val_dtype, val_dtype)
        if code not in _NAMES:
            raise ValueError("field type %r not in %r" % (val_dtype, sorted(_NAMES)))
        self._name, self._val_dtype = name, code

    name = property(lambda self: self._name)
    val_dtype = property(lambda self: self._val_dtype)

    def dtype(self, byte_order="="):
        return byte_order + self._val_dtype

    def __str__(self):
        return "property %s %s" % (_NAMES[self._val_dtype], self._name)

    def __repr__(self):
        return "PlyProperty(%r, %r)" % (self._name, _NAMES[self._val_dtype])


class PlyListProperty(PlyProperty):
    """A variable-length property: a count of type `len_dtype` followed by that many `val_dtype` values per row."""

    def __init__(self, name, len_dtype, val_dtype):
        super().__init__(name, val_dtype)
        code = _TYPES.get(len_dtype, len_dtype)
        if code not in _NAMES or code[0] == "f":
            raise ValueError("list length type %r must be an integer type" % (len_dtype,))
        self._len_dtype = code

    len_dtype = property(lambda self: self._len_dtype)

    def dtype(self, byte_order="="):
        return "|O"

    def list_dtype(self, byte_order="="):
        return np.dtype(byte_order + self._len_dtype), np.dtype(byte_order + self._val_dtype)

    def __str__(self):
        return "property list %s %s %s" % (_NAMES[self._len_dtype], _NAMES[self._val_dtype], self._name)

    def __repr__(self):
        return "PlyListProperty(%r, %r, %r)" % (self._name, _NAMES[self._len_dtype], _NAMES[self._val_dtype])


class PlyElement:
    """A named table: `.name`, `.count`, `.properties`, `.data` (1-D structured numpy array); `el['x']` is a column."""

    def __init__(self, name, properties, count, comments=()):
        name = str(name)
        if not name or any(c.isspace() for c in name):
            raise ValueError("invalid element name %r" % name)
        self._name, self._count = name, int(count)
        self._properties = tuple(properties)
        self._by_name = {p.name: p for p in self._properties}
        if len(self._by_name) != len(self._properties):
            raise ValueError("two properties with same name")
        self.comments = list(comments)
        self._data = None

    name = property(lambda self: self._name)
    count = property(lambda self: self._count)
    properties = property(lambda self: self._properties)

    @property
    def data(self):
        return self._data

    @data.setter
    def data(self, data):
        for p in self._properties:
            if p.name not in data.dtype.fields:
                raise ValueError("dangling property %r" % p.name)
        self._data, self._count = data, len(data)

    def ply_property(self, name):
        return self._by_name[name]

    def dtype(self, byte_order="="):
        return np.dtype([(p.name, p.dtype(byte_order)) for p in self._properties])

    @staticmethod
    def describe(data, name, len_types=None, val_types=None, comments=()):
        """Element over an existing structured array (`scene/gaussian_model.py:207`): one property per field."""
        if not isinstance(data, np.ndarray):
            raise TypeError("only numpy arrays are supported")
        if data.ndim != 1:
            raise ValueError("only one-dimensional arrays are supported")
        if data.dtype.names is None:
            raise ValueError("only structured arrays are supported")
        props = []
        for field in data.dtype.descr:
            if not isinstance(field[1], str):
                raise ValueError("nested records not supported")
            if not field[0]:
                raise ValueError("field with empty name")
            if len(field) != 2:
                raise NotImplementedError("multi-valued field %r: use an object column of arrays" % (field[0],))
            if field[1][1] == "O":
                props.append(PlyListProperty(field[0], (len_types or {}).get(field[0], "u1"),
                                             (val_types or {}).get(field[0], "i4")))
                continue
            code = field[1][1:]
            if code not in _NAMES:
                raise ValueError("unsupported field type %r of %r" % (field[1], field[0]))
            props.append(PlyProperty(field[0], code))
        el = PlyElement(name, props, len(data), comments)
        el.data = data
        return el

    @property
    def header(self):
        lines = ["element %s %d" % (self._name, self._count)]
        lines += ["comment " + c for c in self.comments]
        lines += [str(p) for p in self._properties]
        return "\n".join(lines)

    def __getitem__(self, key):
        return self._data[key]

    def __setitem__(self, key, value):
        self._data[key] = value

    def __len__(self):
        return self._count

    def __contains__(self, name):
        return name in self._by_name

    def __str__(self):
        return self.header

    def __repr__(self):
        return "PlyElement(%r, %r, count=%d, comments=%r)" % (self._name, self._properties, self._count, self.comments)

    # --- body I/O ---------------------------------------------------------------------------------------------------
    def _read(self, stream, text, byte_order):
        dt = self.dtype(byte_order)
        if any(isinstance(p, PlyListProperty) for p in self._properties):
            self._data = self._read_rows(stream, text, byte_order, dt)
        elif text:
            self._data = self._read_txt(stream, dt)
        else:
            need = dt.itemsize * self._count
            raw = stream.read(need)
            if len(raw) < need:
                raise PlyElementParseError("early end-of-file", self, len(raw) // max(dt.itemsize, 1))
            self._data = np.frombuffer(raw, dtype=dt, count=self._count).copy()

    def _read_txt(self, stream, dt):
        out = np.empty(self._count, dtype=dt)
        names = dt.names
        ints = [dt[n].kind in "iu" for n in names]
        for row in range(self._count):
            line = stream.readline()
            if not line:
                raise PlyElementParseError("early end-of-file", self, row)
            fields = line.decode("ascii").split()
            if len(fields) < len(names):
                raise PlyElementParseError("early end-of-line", self, row)
            if len(fields) > len(names):
                raise PlyElementParseError("expected end-of-line", self, row)
            try:
                out[row] = tuple(int(f) if i else float(f) for f, i in zip(fields, ints))
            except ValueError:
                raise PlyElementParseError("malformed input", self, row)
        return out

    def _read_rows(self, stream, text, byte_order, dt):
        """Row-by-row reader for elements with list properties."""
        out = np.empty(self._count, dtype=dt)
        for row in range(self._count):
            if text:
                line = stream.readline()
                if not line:
                    raise PlyElementParseError("early end-of-file", self, row)
                fields = iter(line.decode("ascii").split())
            for p in self._properties:
                try:
                    if isinstance(p, PlyListProperty):
                        lt, vt = p.list_dtype(byte_order)
                        if text:
                            n = int(next(fields))
                            out[p.name][row] = np.array([next(fields) for _ in range(n)], dtype=np.float64).astype(vt)
                        else:
                            raw = stream.read(lt.itemsize)
                            if len(raw) < lt.itemsize:
                                raise PlyElementParseError("early end-of-file", self, row, p)
                            n = int(np.frombuffer(raw, lt)[0])
                            raw = stream.read(n * vt.itemsize)
                            if len(raw) < n * vt.itemsize:
                                raise PlyElementParseError("early end-of-file", self, row, p)
                            out[p.name][row] = np.frombuffer(raw, vt).astype(vt.newbyteorder("="))
                    elif text:
                        f = next(fields)
                        out[p.name][row] = int(f) if p.val_dtype[0] in "iu" else float(f)
                    else:
                        vt = np.dtype(p.dtype(byte_order))
                        raw = stream.read(vt.itemsize)
                        if len(raw) < vt.itemsize:
                            raise PlyElementParseError("early end-of-file", self, row, p)
                        out[p.name][row] = np.frombuffer(raw, vt)[0]
                except StopIteration:
                    raise PlyElementParseError("early end-of-line", self, row, p)
                except ValueError:
                    raise PlyElementParseError("malformed input", self, row, p)
            if text and next(fields, None) is not None:
                raise PlyElementParseError("expected end-of-line", self, row)
        return out

    def _write_rows(self, stream, text, byte_order):
        for rec in self._data:
            parts = []
            for p in self._properties:
                v = rec[p.name]
                if isinstance(p, PlyListProperty):
                    lt, vt = p.list_dtype(byte_order)
                    v = np.asarray(v)
                    if text:
                        parts.append("%d" % v.size)
                        parts += [repr(float(x)) if vt.kind == "f" else "%d" % x for x in v.ravel()]
                    else:
                        parts.append(np.array([v.size], lt).tobytes() + v.astype(vt).tobytes())
                elif text:
                    parts.append(repr(float(v)) if p.val_dtype[0] == "f" else "%d" % v)
                else:
                    parts.append(np.array([v], np.dtype(p.dtype(byte_order))).tobytes())
            stream.write((" ".join(parts) + "\n").encode("ascii") if text else b"".join(parts))

    def _write(self, stream, text, byte_order):
        if any(isinstance(p, PlyListProperty) for p in self._properties):
            self._write_rows(stream, text, byte_order)
        elif text:
            for rec in self._data:
                fields = []
                for p in self._properties:
                    v = rec[p.name]
                    fields.append(repr(float(v)) if p.val_dtype[0] == "f" else "%d" % v)
                stream.write((" ".join(fields) + "\n").encode("ascii"))
        else:
            # Field by field into a packed record of the file's byte order (the array may be padded or foreign-endian).
            packed = np.empty(self._count, dtype=self.dtype(byte_order))
            for p in self._properties:
                packed[p.name] = self._data[p.name]
            stream.write(packed.tobytes())


class PlyData:
    """A PLY file: `.elements`, `.text`, `.byte_order`, `.comments`, `.obj_info`; `ply['vertex']` finds by name."""

    def __init__(self, elements=(), text=False, byte_order="=", comments=(), obj_info=()):
        if byte_order == "=" and not text:
            byte_order = _NATIVE
        if byte_order not in ("<", ">", "="):
            raise ValueError("byte order must be '<', '>', or '='")
        self.byte_order, self.text = byte_order, bool(text)
        self.comments, self.obj_info = list(comments), list(obj_info)
        self.elements = elements

    @property
    def elements(self):
        return self._elements

    @elements.setter
    def elements(self, elements):
        self._elements = tuple(elements)
        self._by_name = {e.name: e for e in self._elements}
        if len(self._by_name) != len(self._elements):
            raise ValueError("two elements with same name")

    def __getitem__(self, name):
        return self._by_name[name]

    def __contains__(self, name):
        return name in self._by_name

    def __iter__(self):
        return iter(self._elements)

    def __len__(self):
        return len(self._elements)

    @property
    def header(self):
        fmt = "ascii" if self.text else {"<": "binary_little_endian", ">": "binary_big_endian"}[self.byte_order]
        lines = ["ply", "format %s 1.0" % fmt]
        lines += ["comment " + c for c in self.comments]
        lines += ["obj_info " + c for c in self.obj_info]
        lines += [e.header for e in self._elements]
        lines.append("end_header")
        return "\n".join(lines)

    def __str__(self):
        return self.header

    # --- reading ----------------------------------------------------------------------------------------------------
    @staticmethod
    def _parse_header(stream):
        def next_line(n):
            raw = stream.readline()
            if not raw:
                raise PlyHeaderParseError("early end-of-file", n)
            try:
                return raw.decode("ascii").strip()
            except UnicodeDecodeError:
                raise PlyHeaderParseError("non-ASCII header", n)

        n = 1
        if next_line(n) != "ply":
            raise PlyHeaderParseError("expected 'ply'", n)
        n += 1
        fields = next_line(n).split()
        if len(fields) != 3 or fields[0] != "format":
            raise PlyHeaderParseError("expected 'format'", n)
        if fields[1] not in _FORMATS:
            raise PlyHeaderParseError("don't understand format %r" % fields[1], n)
        if fields[2] != "1.0":
            raise PlyHeaderParseError("expected version '1.0'", n)
        fmt = fields[1]
        comments, obj_info, specs = [], [], []   # specs: [name, count, [PlyProperty], [comments]]
        while True:
            n += 1
            line = next_line(n)
            key, _, rest = line.partition(" ")
            if key == "end_header":
                break
            if key == "comment":
                (specs[-1][3] if specs else comments).append(rest)
            elif key == "obj_info":
                obj_info.append(rest)
            elif key == "element":
                f = rest.split()
                if len(f) != 2:
                    raise PlyHeaderParseError("expected 'element <name> <count>'", n)
                try:
                    count = int(f[1])
                except ValueError:
                    raise PlyHeaderParseError("expected integer", n)
                if count < 0:
                    raise PlyHeaderParseError("negative element count", n)
                specs.append([f[0], count, [], []])
            elif key == "property":
                if not specs:
                    raise PlyHeaderParseError("property without element", n)
                f = rest.split()
                if f and f[0] == "list":
                    if len(f) != 4:
                        raise PlyHeaderParseError("expected 'property list <len type> <type> <name>'", n)
                    if f[1] not in _TYPES or f[2] not in _TYPES:
                        raise PlyHeaderParseError("unknown data type in %r" % rest, n)
                    try:
                        specs[-1][2].append(PlyListProperty(f[3], f[1], f[2]))
                    except ValueError as e:
                        raise PlyHeaderParseError(str(e), n)
                    continue
                if len(f) != 2:
                    raise PlyHeaderParseError("expected 'property <type> <name>'", n)
                if f[0] not in _TYPES:
                    raise PlyHeaderParseError("unknown data type %r" % f[0], n)
                specs[-1][2].append(PlyProperty(f[1], f[0]))
            elif not line:
                continue
            else:
                raise PlyHeaderParseError("unexpected keyword %r" % key, n)
        elements = [PlyElement(name, props, count, cm) for name, count, props, cm in specs]
        return fmt, comments, obj_info, elements

    @staticmethod
    def read(stream, mmap=False, known_list_len=None):
        """Reads a whole file from a path or a binary stream (`scene/gaussian_model.py:216`)."""
        own = isinstance(stream, (str, bytes)) or hasattr(stream, "__fspath__")
        f = open(stream, "rb") if own else stream
        try:
            fmt, comments, obj_info, elements = PlyData._parse_header(f)
            ply = PlyData(elements, text=(fmt == "ascii"), byte_order=_FORMATS[fmt], comments=comments,
                          obj_info=obj_info)
            for el in ply.elements:
                el._read(f, ply.text, ply.byte_order)
        finally:
            if own:
                f.close()
        return ply

    # --- writing ----------------------------------------------------------------------------------------------------
    def write(self, stream):
        """Writes header + elements to a path or a binary stream (`scene/gaussian_model.py:208`)."""
        own = isinstance(stream, (str, bytes)) or hasattr(stream, "__fspath__")
        f = open(stream, "wb") if own else stream
        try:
            f.write((self.header + "\n").encode("ascii"))
            for el in self._elements:
                el._write(f, self.text, self.byte_order)
        finally:
            if own:
                f.close()
