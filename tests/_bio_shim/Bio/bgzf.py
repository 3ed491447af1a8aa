"""TEST-ONLY stand-in for the part of Bio.bgzf that kmerdb/fileutil.py builds on (KDBReader subclasses BgzfReader and
calls _load_block / readline / BgzfBlocks; KDBWriter subclasses BgzfWriter).  Behaviour follows the BGZF format
(gzip members with a 'BC' extra subfield holding the member size, <= 64 KiB of payload each) and Biopython 1.83's
documented reader semantics, written from the format description — Biopython itself is absent from this image.
Used only by tests/golden/make_golden_kdb.py to run the UNMODIFIED reference reader/writer."""
import struct
import zlib

_MAGIC = b"\x1f\x8b\x08\x04"
_EOF = bytes.fromhex("1f8b08040000000000ff0600424302001b0003000000000000000000")


def _load_bgzf_block(handle, text_mode=False):
    magic = handle.read(4)
    if not magic:
        raise StopIteration
    if magic != _MAGIC:
        raise ValueError("A BGZF block should start with %r, not %r" % (_MAGIC, magic))
    _mtime, _xfl, _os, extra_len = struct.unpack("<LBBH", handle.read(8))
    block_size, x = None, 0
    while x < extra_len:
        sid = handle.read(2)
        slen = struct.unpack("<H", handle.read(2))[0]
        sdata = handle.read(slen)
        x += slen + 4
        if sid == b"BC":
            block_size = struct.unpack("<H", sdata)[0] + 1
    if block_size is None:
        raise ValueError("Missing BC, this isn't a BGZF file!")
    d = zlib.decompressobj(-15)
    data = d.decompress(handle.read(block_size - 1 - extra_len - 19)) + d.flush()
    crc = struct.unpack("<I", handle.read(4))[0]
    size = struct.unpack("<I", handle.read(4))[0]
    if size != len(data) or crc != (zlib.crc32(data) & 0xFFFFFFFF):
        raise RuntimeError("BGZF block fails its CRC / length check")
    return block_size, (data.decode("latin-1") if text_mode else data)


def BgzfBlocks(handle):
    """(raw start, raw length, data start, data length) of every block from the handle's position on."""
    data_start = 0
    while True:
        start = handle.tell()
        try:
            block_length, data = _load_bgzf_block(handle)
        except StopIteration:
            break
        yield start, block_length, data_start, len(data)
        data_start += len(data)


class BgzfReader:
    def __init__(self, filename=None, mode="r", fileobj=None, max_cache=100):
        handle = fileobj if fileobj else open(filename, "rb")
        self._text = "b" not in mode.lower()
        self._newline = "\n" if self._text else b"\n"
        self._handle = handle
        self.max_cache = max_cache
        self._buffers = {}
        self._block_start_offset = None
        self._block_raw_length = None
        self._load_block(handle.tell())

    def _load_block(self, start_offset=None):
        if start_offset is None:
            start_offset = self._block_start_offset + self._block_raw_length
        if start_offset == self._block_start_offset:
            self._within_block_offset = 0
            return
        if start_offset in self._buffers:
            self._buffer, self._block_raw_length = self._buffers[start_offset]
            self._within_block_offset = 0
            self._block_start_offset = start_offset
            return
        while len(self._buffers) >= self.max_cache:
            self._buffers.popitem()
        handle = self._handle
        if start_offset != handle.tell():
            handle.seek(start_offset)
        self._block_start_offset = handle.tell()
        try:
            block_size, self._buffer = _load_bgzf_block(handle, self._text)
        except StopIteration:
            block_size, self._buffer = 0, ("" if self._text else b"")
        self._within_block_offset = 0
        self._block_raw_length = block_size
        self._buffers[self._block_start_offset] = self._buffer, block_size

    def tell(self):
        if 0 < self._within_block_offset and self._within_block_offset == len(self._buffer):
            return (self._block_start_offset + self._block_raw_length) << 16
        return (self._block_start_offset << 16) | self._within_block_offset

    def seek(self, virtual_offset):
        start_offset, within_block = virtual_offset >> 16, virtual_offset & 0xFFFF
        if start_offset != self._block_start_offset:
            self._load_block(start_offset)
        if within_block > len(self._buffer) and not (within_block == 0 and len(self._buffer) == 0):
            raise ValueError("Within offset %i but block size only %i" % (within_block, len(self._buffer)))
        self._within_block_offset = within_block
        return virtual_offset

    def read(self, size=-1):
        result = "" if self._text else b""
        while size and self._block_raw_length:
            rest = len(self._buffer) - self._within_block_offset
            if 0 <= size < rest:
                result += self._buffer[self._within_block_offset:self._within_block_offset + size]
                self._within_block_offset += size
                break
            result += self._buffer[self._within_block_offset:]
            size = size - rest if size > 0 else size
            self._load_block()
        return result

    def readline(self):
        result = "" if self._text else b""
        while self._block_raw_length:
            i = self._buffer.find(self._newline, self._within_block_offset)
            if i == -1:
                result += self._buffer[self._within_block_offset:]
                self._load_block()
            elif i + 1 == len(self._buffer):
                result += self._buffer[self._within_block_offset:]
                self._load_block()
                break
            else:
                result += self._buffer[self._within_block_offset:i + 1]
                self._within_block_offset = i + 1
                break
        return result

    def __next__(self):
        line = self.readline()
        if not line:
            raise StopIteration
        return line

    def __iter__(self):
        return self

    def close(self):
        self._handle.close()
        self._buffer = None
        self._block_start_offset = None
        self._buffers = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


class BgzfWriter:
    def __init__(self, filename=None, mode="w", fileobj=None, compresslevel=6):
        self._handle = fileobj if fileobj else open(filename, "ab" if "a" in mode.lower() else "wb")
        self._text = "b" not in mode.lower()
        self._buffer = b""
        self.compresslevel = compresslevel

    def _write_block(self, block):
        c = zlib.compressobj(self.compresslevel, zlib.DEFLATED, -15, zlib.DEF_MEM_LEVEL, 0)
        compressed = c.compress(block) + c.flush()
        if len(compressed) > 65536 - 26:
            c = zlib.compressobj(0, zlib.DEFLATED, -15, zlib.DEF_MEM_LEVEL, 0)
            compressed = c.compress(block) + c.flush()
        bsize = struct.pack("<H", len(compressed) + 25)
        crc = struct.pack("<I", zlib.crc32(block) & 0xFFFFFFFF)
        self._handle.write(_MAGIC + b"\x00\x00\x00\x00\x00\xff\x06\x00\x42\x43\x02\x00" + bsize + compressed + crc + struct.pack("<I", len(block)))

    def write(self, data):
        if isinstance(data, str):
            data = data.encode("latin-1")
        self._buffer += data
        while len(self._buffer) >= 65536:
            self._write_block(self._buffer[:65536])
            self._buffer = self._buffer[65536:]

    def flush(self):
        while len(self._buffer) >= 65536:
            self._write_block(self._buffer[:65535])
            self._buffer = self._buffer[65535:]
        self._write_block(self._buffer)
        self._buffer = b""
        self._handle.flush()

    def close(self):
        if self._buffer:
            self.flush()
        self._handle.write(_EOF)
        self._handle.flush()
        self._handle.close()

    def tell(self):
        return self._handle.tell() << 16

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
