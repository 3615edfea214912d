"""CPU: the bit buffer / Huffman / `.fqt` + HV file formats stay byte-compatible with the reference.

Known answers: the two bundled `.fqt` files (tests/golden, copies of the reference's imgs/compressed), the reference's
own self-check vectors for BitBuffer (bitbuffer.py:70-105) and Huffman (huffman.py:124-157) as observed by the survey.
"""
import io
import struct

import numpy as np
import pytest

from test_oracle_golden import hv_params, qt_params


def qt_image_from_golden(g, name, shape):
    from quadtree.common import EncodingInfo, QuadtreeImage, stream_to_forest
    p = qt_params(g, name)
    nodes = np.c_[g[f"{name}_nodes"], np.zeros(len(g[f"{name}_nodes"]), np.int32)]
    vals = np.c_[g[f"{name}_vals"], np.zeros(len(g[f"{name}_vals"]))]
    forest = stream_to_forest(nodes, vals, int(g[f"{name}_nroots"]))
    return QuadtreeImage(EncodingInfo(p[0], p[1], p[2], shape[1], shape[0]), forest)


def test_bitbuffer_known_answers():
    from bitbuffer import BitBuffer
    out = io.BytesIO()
    bb = BitBuffer(out)
    emitted = [bb.write(346, 16), bb.write(200, 8), bb.write(3434, 16), bb.write(1, 24), bb.write(1, 32),
               bb.write(0x4FF, 11), bb.write(0xFFFFF, 21), bb.write(0xFFFFFFFF, 32)]
    words = [e.hex() for e in emitted if e is not None]
    assert words == ["015ac80d", "6a000001", "00000001", "9fefffff", "ffffffff"]
    src = io.BytesIO(struct.pack('!III', 0b0110101011, 0b0111 << 28, 12))
    bb = BitBuffer(src)
    got = [bb.read(22), bb.read(1), bb.read(1), bb.read(1), bb.read(1)]
    bb.read(5)
    got.append(bb.read(3))
    assert got == [0, 0, 1, 1, 0, 5]
    assert BitBuffer(io.BytesIO(b"\x00\x01")).read(3) == -1                         # EOF on a short tail
    out = io.BytesIO(); bb = BitBuffer(out); bb.write(5, 3); assert not bb.is_empty(); bb.flush()
    assert out.getvalue() == bytes([0b10100000, 0, 0, 0]) and bb.is_empty()


def test_huffman_round_trip_and_tree_serialisation():
    from bitbuffer import BitBuffer
    from entropy.huffman import HuffmanDecoder, HuffmanDeserializer, HuffmanEncoder, HuffmanSerializer
    symbols = [5, 5, 5, 5, 5, 1, 2, 3, 1, 1, 2, 2, 3]
    enc = HuffmanEncoder()
    table = enc.prepare_for_encoding(symbols)
    # known answer produced by the reference's HuffmanEncoder on its own self-check input (huffman.py:125)
    assert table == {3: (0, 2), 2: (1, 2), 1: (2, 2), 5: (3, 2)}
    buf = io.BytesIO()
    bb = BitBuffer(buf)
    for s in symbols:
        bb.write(*enc.encode(s))
    bb.flush()
    buf.seek(0)
    dec = HuffmanDecoder(enc.root)
    rb = BitBuffer(buf)
    assert [dec.decode(rb) for _ in symbols] == symbols
    buf = io.BytesIO()
    bb = BitBuffer(buf)
    HuffmanSerializer().serialize(enc.root, 8, bb)
    bb.flush()
    buf.seek(0)
    root = HuffmanDeserializer().deserialize(BitBuffer(buf), 8)
    assert HuffmanEncoder()._build_coding_table(root) == table
    single = HuffmanEncoder()
    assert single.prepare_for_encoding([7, 7, 7]) == {7: (0, 0)}                    # one symbol: empty code


@pytest.mark.parametrize("name,fqt", [("cauliflower256", "kalafior256.fqt"), ("lena313x413", "lena313x413.fqt")])
def test_fqt_files_are_reproduced_byte_for_byte(golden, images, name, fqt):
    from entropy.processor import CodingDimension, HuffmanEntropyProcessor
    from quadtree.serialization import QuadtreeDeserializer, QuadtreeSerializer
    want = open(golden.path(fqt), "rb").read()
    q = qt_image_from_golden(golden["qt_golden"], name, images[name].shape)
    out = io.BytesIO()
    QuadtreeSerializer(HuffmanEntropyProcessor(), set(CodingDimension)).serialize(q, out)
    assert out.getvalue() == want
    back = QuadtreeDeserializer().deserialize(io.BytesIO(want))
    assert (back.info.img_width, back.info.img_height, back.info.scale_bits, back.info.offset_bits) == \
        (q.info.img_width, q.info.img_height, 7, 5)
    again = io.BytesIO()
    QuadtreeSerializer(HuffmanEntropyProcessor(), set(CodingDimension)).serialize(back, again)
    assert again.getvalue() == want
    plain = io.BytesIO()
    QuadtreeSerializer(HuffmanEntropyProcessor(), set()).serialize(q, plain)
    if name == "cauliflower256":
        assert len(plain.getvalue()) == 16566


@pytest.mark.parametrize("name,fqt", [("cauliflower256", "kalafior256.fqt"), ("lena313x413", "lena313x413.fqt"),
                                      ("cauliflower240x92", None), ("cauliflower64", None)])
def test_flat_stream_paths_equal_object_paths(orc, golden, images, name, fqt):
    """An image that still carries the encoder's flat stream is flattened for the decoder and written to a file with
    array operations; both must give exactly what walking the node objects gives (and the bundled files)."""
    from entropy.processor import CodingDimension, HuffmanEntropyProcessor
    from quadtree.common import EncodingInfo, QuadtreeImage, forest_to_leaves, stream_to_forest
    from quadtree.serialization import QuadtreeSerializer
    p = qt_params(golden["qt_golden"], name)
    img = images[name]
    nodes, vals, n_roots = orc.qt_encode(img, *p)           # the stream layout of fic_qt_encode, exponent column included
    info = EncodingInfo(p[0], p[1], p[2], img.shape[1], img.shape[0])

    def lazy():
        return QuadtreeImage(info, stream=(nodes, vals, n_roots))

    objects = QuadtreeImage(info, stream_to_forest(nodes, vals, n_roots))
    assert objects.flat_stream() is None
    for quantised in (True, False):
        a, b = forest_to_leaves(lazy(), quantised), forest_to_leaves(objects, quantised)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    for dims in (set(CodingDimension), set(), {CodingDimension.SCALE}):
        fast, slow = io.BytesIO(), io.BytesIO()
        q = lazy()
        QuadtreeSerializer(HuffmanEntropyProcessor(), set(dims)).serialize(q, fast)
        assert q.flat_stream() is not None                  # writing a file does not build the objects
        QuadtreeSerializer(HuffmanEntropyProcessor(), set(dims)).serialize(objects, slow)
        assert fast.getvalue() == slow.getvalue()
        if fqt and dims == set(CodingDimension):
            assert fast.getvalue() == open(golden.path(fqt), "rb").read()
    q = lazy()
    assert len(q.forest) == n_roots and q.flat_stream() is None      # first look at the objects drops the stream
    q.forest = objects.forest
    assert q == objects


def _preorder(node, out, children):
    out.append(node)
    for c in children(node):
        _preorder(c, out, children)
    return out


@pytest.mark.parametrize("fqt", ["kalafior256.fqt", "lena313x413.fqt"])
def test_native_fqt_reader_equals_bit_by_bit_reader(golden, fqt):
    """`fic_tree_read_nodes` (the reader the deserialiser uses) against the field-by-field Python loop on the bundled files."""
    from _stream import read_preamble
    from bitbuffer import BitBuffer
    from quadtree.serialization import QUADTREE_MAGIC, QuadtreeDeserializer
    blob = open(golden.path(fqt), "rb").read()
    fast = QuadtreeDeserializer().deserialize(io.BytesIO(blob))
    assert fast.flat_stream() is not None
    f = io.BytesIO(blob)
    bits = BitBuffer(f, "!")
    slow = QuadtreeDeserializer()
    info, slow.size, slow.entropy_processor = read_preamble(f, bits, QUADTREE_MAGIC, 'quadtree')
    forest = slow._read_forest(bits, info)
    assert info == fast.info and len(forest) == len(fast.forest)
    for a, b in zip(forest, fast.forest):
        na = _preorder(a, [], lambda n: n.children or [])
        nb = _preorder(b, [], lambda n: n.children or [])
        assert len(na) == len(nb)
        assert all((x.domain is None) == (y.domain is None) and x.domain == y.domain for x, y in zip(na, nb))


def test_native_reader_reports_truncated_files(golden):
    from quadtree.serialization import QuadtreeDeserializer
    blob = open(golden.path("kalafior256.fqt"), "rb").read()
    with pytest.raises(AssertionError, match="EOF"):        # the reference asserts (serialization.py:188)
        QuadtreeDeserializer().deserialize(io.BytesIO(blob[:len(blob) // 2]))


@pytest.mark.parametrize("name", ["lena64", "lena48x80", "cauliflower240x92", "lena256"])
def test_hv_flat_stream_paths_equal_object_paths(golden, name):
    """HV images that still carry their flat stream: leaf table, file writer and native reader against the object walks."""
    from _stream import read_preamble
    from bitbuffer import BitBuffer
    from entropy.processor import CodingDimension, HuffmanEntropyProcessor
    from hvv2.common import HVImage, stream_to_tree, tree_to_leaves
    from hvv2.serialization import HV_MAGIC, HVDeserializer, HVSerializer
    from quadtree.common import EncodingInfo
    g = golden["hv_golden"]
    p = hv_params(g, name)
    H, W = g[f"{name}_dec"].shape
    nodes = np.ascontiguousarray(np.c_[g[f"{name}_nodes"], np.zeros(len(g[f"{name}_nodes"]), np.int32)], np.int32)
    vals = np.c_[g[f"{name}_vals"], np.zeros(len(g[f"{name}_vals"]))]
    info = EncodingInfo(p[4], p[5], 1.0, W, H)
    objects = HVImage(info, stream_to_tree(nodes, vals))
    assert objects.flat_stream() is None
    for quantised in (True, False):
        a, b = tree_to_leaves(HVImage(info, stream=(nodes, vals)), quantised), tree_to_leaves(objects, quantised)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    blobs = []
    for dims in (set(CodingDimension), set()):
        fast, slow = io.BytesIO(), io.BytesIO()
        lazy = HVImage(info, stream=(nodes, vals))
        HVSerializer(HuffmanEntropyProcessor(), set(dims)).serialize(lazy, fast)
        assert lazy.flat_stream() is not None
        HVSerializer(HuffmanEntropyProcessor(), set(dims)).serialize(objects, slow)
        assert fast.getvalue() == slow.getvalue()
        blobs.append(fast.getvalue())
    for blob in blobs:
        back = HVDeserializer().deserialize(io.BytesIO(blob))
        got = back.flat_stream()
        assert got is not None and np.array_equal(got[0][:, :7], nodes[:, :7])
        f = io.BytesIO(blob)
        bits = BitBuffer(f, "!")
        slow = HVDeserializer()
        _, slow.size, slow.entropy_processor = read_preamble(f, bits, HV_MAGIC, 'hv')
        a, b = tree_to_leaves(HVImage(info, slow._read_tree(bits))), tree_to_leaves(back)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_legacy_fqt_header_is_rejected_like_the_reference():
    from quadtree.serialization import QuadtreeDeserializer
    legacy = struct.pack("!8sfBBHHI", b"quadtree", 1.0, 7, 5, 380, 322, 0) + b"\x00" * 8   # entropy id 0
    with pytest.raises(ValueError):
        QuadtreeDeserializer().deserialize(io.BytesIO(legacy))
    with pytest.raises(Exception, match="magic"):
        QuadtreeDeserializer().deserialize(io.BytesIO(b"notquadt" + b"\x00" * 30))
    with pytest.raises(Exception, match="too short"):
        QuadtreeDeserializer().deserialize(io.BytesIO(b"abc"))


@pytest.mark.parametrize("name", ["lena64", "cauliflower240x92"])
def test_hv_stream_round_trip(golden, name):
    from entropy.processor import CodingDimension, HuffmanEntropyProcessor
    from hvv2.common import HVImage, stream_to_tree
    from hvv2.serialization import HVDeserializer, HVSerializer
    from quadtree.common import EncodingInfo
    g = golden["hv_golden"]
    p = hv_params(g, name)
    img = g[f"{name}_img"]
    nodes = np.c_[g[f"{name}_nodes"], np.zeros(len(g[f"{name}_nodes"]), np.int32)]
    vals = np.c_[g[f"{name}_vals"], np.zeros(len(g[f"{name}_vals"]))]
    hv = HVImage(EncodingInfo(p[4], p[5], 1.0, img.shape[1], img.shape[0]), stream_to_tree(nodes, vals))
    out = io.BytesIO()
    HVSerializer(HuffmanEntropyProcessor(), set(CodingDimension)).serialize(hv, out)
    data = out.getvalue()
    assert data[:2] == b"hv" and struct.unpack("!I", data[12:16])[0] == len(data) and len(data) % 4 == 0
    back = HVDeserializer().deserialize(io.BytesIO(data))

    def flat(root):
        rows, stack = [], [root]
        while stack:
            n = stack.pop()
            if n.is_leaf():
                d = n.domain
                rows.append((1, d.start_i, d.start_j, d.orientation, d.rotation, d.quantized_scale, d.quantized_offset))
            else:
                rows.append((0, int(n.split_info.vertical_split), n.split_info.split_idx, 0, 0, 0, 0))
                stack += [n.c2, n.c1]
        return rows
    assert flat(back.root) == flat(hv.root)
    again = io.BytesIO()
    HVSerializer(HuffmanEntropyProcessor(), set(CodingDimension)).serialize(back, again)
    assert again.getvalue() == data


def test_rgb_stream_is_three_fqt_streams(golden, images):
    from entropy.processor import CodingDimension, HuffmanEntropyProcessor
    from quadtree.common import ColoredQuadtreeImage
    from quadtree.rgb import RGBQuadtreeDeserializer, RGBQuadtreeSerializer
    from quadtree.serialization import QuadtreeDeserializer, QuadtreeSerializer
    g = golden["qt_golden"]
    q1 = qt_image_from_golden(g, "cauliflower64", images["cauliflower64"].shape)
    q2 = qt_image_from_golden(g, "cauliflower64step4", images["cauliflower64"].shape)
    out = io.BytesIO()
    # planes of equal size (4:4:4): like the reference, one serialiser object keeps its field widths and coders from
    # the first plane (entropy/processor.py:88-98), so differently sized planes need separate serialisers
    RGBQuadtreeSerializer(QuadtreeSerializer(HuffmanEntropyProcessor(), set())).serialize(ColoredQuadtreeImage(q1, q2, q1), out)
    data = out.getvalue()
    back = RGBQuadtreeDeserializer(QuadtreeDeserializer()).deserialize(io.BytesIO(data))
    one = io.BytesIO()
    QuadtreeSerializer(HuffmanEntropyProcessor(), set()).serialize(q1, one)
    assert data.startswith(one.getvalue()) and data.endswith(one.getvalue())
    for plane, src in ((back.encoded_Y, q1), (back.encoded_Cb, q2), (back.encoded_Cr, q1)):
        again = io.BytesIO()
        QuadtreeSerializer(HuffmanEntropyProcessor(), set()).serialize(plane, again)
        ref = io.BytesIO()
        QuadtreeSerializer(HuffmanEntropyProcessor(), set()).serialize(src, ref)
        assert again.getvalue() == ref.getvalue()


def _random_quadtree_stream(rng, width, height, min_side=2, p_split=0.6):
    """A random forest over a (possibly ragged) image as the preorder stream of fic_qt_encode."""
    from quadtree.common import iter_root_squares
    nodes = []

    def grow(e):
        side = 1 << e
        if side > min_side and (2 * side > min(width, height) or rng.random() < p_split):   # a leaf's 2x domain must fit
            nodes.append([0, 0, 0, 0, 0, 0, 0, e])
            for _ in range(4):
                grow(e - 1)
        else:
            nodes.append([1, int(rng.integers(0, max(1, height - 2 * side + 1))), int(rng.integers(0, max(1, width - 2 * side + 1))),
                          int(rng.integers(0, 2)), int(rng.integers(0, 4)), int(rng.integers(0, 128)), int(rng.integers(0, 32)), e])

    roots = list(iter_root_squares(width, height))
    for _, _, e in roots:
        grow(e)
    nodes = np.array(nodes, np.int32)
    vals = np.c_[rng.normal(size=(len(nodes), 2)), np.zeros(len(nodes))]
    vals[nodes[:, 0] == 0] = 0.0
    return nodes, vals, len(roots)


@pytest.mark.parametrize("seed,shape", [(1, (64, 64)), (2, (70, 45)), (3, (33, 96)), (4, (128, 40)), (5, (17, 23)), (6, (256, 256))])
def test_random_forests_flat_paths_equal_object_paths_and_round_trip(seed, shape):
    """Random quadtree forests (ragged images, uneven depths): array-based flattening, file writer and native reader
    against the object walks, and writer -> reader -> writer reproduces the bytes."""
    from entropy.processor import CodingDimension, HuffmanEntropyProcessor
    from quadtree.common import EncodingInfo, QuadtreeImage, forest_to_leaves, stream_to_forest
    from quadtree.serialization import QuadtreeDeserializer, QuadtreeSerializer
    rng = np.random.default_rng(seed)
    H, W = shape
    nodes, vals, n_roots = _random_quadtree_stream(rng, W, H)
    info = EncodingInfo(7, 5, 1.0, W, H)
    objects = QuadtreeImage(info, stream_to_forest(nodes, vals, n_roots))
    for quantised in (True, False):
        a = forest_to_leaves(QuadtreeImage(info, stream=(nodes, vals, n_roots)), quantised)
        b = forest_to_leaves(objects, quantised)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    leaves = forest_to_leaves(objects)[0]
    cover = np.zeros((H, W), np.int32)
    for ri, rj, hh, ww in leaves[:, :4]:
        cover[ri:ri + hh, rj:rj + ww] += 1
    assert (cover == 1).all()                                   # the leaves tile the image exactly
    for dims in (set(CodingDimension), {CodingDimension.DOMAIN_ROW, CodingDimension.OFFSET}, set()):
        fast, slow = io.BytesIO(), io.BytesIO()
        QuadtreeSerializer(HuffmanEntropyProcessor(), set(dims)).serialize(QuadtreeImage(info, stream=(nodes, vals, n_roots)), fast)
        QuadtreeSerializer(HuffmanEntropyProcessor(), set(dims)).serialize(objects, slow)
        assert fast.getvalue() == slow.getvalue()
        back = QuadtreeDeserializer().deserialize(io.BytesIO(fast.getvalue()))
        got = back.flat_stream()
        assert got is not None and np.array_equal(got[0], nodes) and got[2] == n_roots
        again = io.BytesIO()
        QuadtreeSerializer(HuffmanEntropyProcessor(), set(dims)).serialize(back, again)
        assert again.getvalue() == fast.getvalue()


def _random_hv_stream(rng, width, height, min_side=2, p_split=0.7):
    nodes = []

    def grow(w, h):
        can_v, can_h = w >= 2 * min_side, h >= 2 * min_side
        too_wide, too_tall = 2 * w > width, 2 * h > height       # a leaf's 2x domain must fit the image
        if (can_v or can_h) and (too_wide or too_tall or rng.random() < p_split):
            vertical = can_v and (not can_h or rng.random() < 0.5)
            if too_wide and can_v:
                vertical = True
            elif too_tall and can_h:
                vertical = False
            size = w if vertical else h
            s = int(rng.integers(min_side - 1, size - min_side))   # first child keeps s + 1 columns / rows
            nodes.append([0, int(vertical), s, 0, 0, 0, 0, 0])
            if vertical:
                grow(s + 1, h); grow(w - s - 1, h)
            else:
                grow(w, s + 1); grow(w, h - s - 1)
        else:
            nodes.append([1, int(rng.integers(0, max(1, height - 2 * h + 1))), int(rng.integers(0, max(1, width - 2 * w + 1))),
                          int(rng.integers(0, 2)), 2 * int(rng.integers(0, 2)), int(rng.integers(0, 32)), int(rng.integers(0, 128)), 0])

    grow(width, height)
    nodes = np.array(nodes, np.int32)
    vals = np.c_[rng.normal(size=(len(nodes), 2)), np.zeros(len(nodes))]
    return nodes, vals


@pytest.mark.parametrize("seed,shape", [(1, (64, 64)), (2, (70, 45)), (3, (33, 96)), (4, (200, 120))])
def test_random_hv_trees_flat_paths_equal_object_paths_and_round_trip(seed, shape):
    from entropy.processor import CodingDimension, HuffmanEntropyProcessor
    from hvv2.common import HVImage, stream_to_tree, tree_to_leaves
    from hvv2.serialization import HVDeserializer, HVSerializer
    from quadtree.common import EncodingInfo
    rng = np.random.default_rng(seed)
    H, W = shape
    nodes, vals = _random_hv_stream(rng, W, H)
    info = EncodingInfo(5, 7, 1.0, W, H)
    objects = HVImage(info, stream_to_tree(nodes, vals))
    for quantised in (True, False):
        a = tree_to_leaves(HVImage(info, stream=(nodes, vals)), quantised)
        b = tree_to_leaves(objects, quantised)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    cover = np.zeros((H, W), np.int32)
    for ri, rj, hh, ww in tree_to_leaves(objects)[0][:, :4]:
        cover[ri:ri + hh, rj:rj + ww] += 1
    assert (cover == 1).all()
    for dims in (set(CodingDimension), set()):
        fast, slow = io.BytesIO(), io.BytesIO()
        HVSerializer(HuffmanEntropyProcessor(), set(dims)).serialize(HVImage(info, stream=(nodes, vals)), fast)
        HVSerializer(HuffmanEntropyProcessor(), set(dims)).serialize(objects, slow)
        assert fast.getvalue() == slow.getvalue()
        back = HVDeserializer().deserialize(io.BytesIO(fast.getvalue()))
        assert np.array_equal(back.flat_stream()[0][:, :7], nodes[:, :7])


def test_bruteforce_encoded_image_builds_domains_on_demand():
    from bruteforce.model import Domain, EncodedImage
    rng = np.random.default_rng(3)
    idx = rng.integers(0, 40, (16, 4)).astype(np.int32)
    so = rng.normal(size=(16, 2))
    enc = EncodedImage.from_arrays(32, 32, 8, idx, so)
    assert len(enc) == 16
    a, b = enc.to_arrays()
    assert np.array_equal(a, idx) and np.array_equal(b, so) and enc._domains is None     # no objects were needed
    doms = enc.domains
    assert enc._arrays is None and len(doms) == 16 and isinstance(doms[3], Domain)
    assert (doms[3].start_i, doms[3].rotation, doms[3].o) == (int(idx[3, 0]), int(idx[3, 3]), float(so[3, 1]))
    doms[0].s = 0.25                                                                        # edits are seen by to_arrays
    assert enc.to_arrays()[1][0, 0] == 0.25
    fresh = EncodedImage(32, 32, 8)
    fresh.domains.append(Domain(1, 2, 0, 3, 0.5, 7.0))                                      # the reference's own usage
    assert fresh.to_arrays()[0].tolist() == [[1, 2, 0, 3]] and len(fresh) == 1
    fresh.domains = []
    assert fresh.to_arrays()[0].shape == (0, 4)


def test_symbol_statistics_and_column_encoding_equal_the_scalar_forms():
    """`ordered_counts` on arrays = `Counter(list).items()` (first-occurrence order decides the Huffman tree shape), and
    `encode_column` = one `encode` per element, for the Huffman and the identity encoder."""
    from collections import Counter
    from entropy.abccodec import IdentityEncoder
    from entropy.huffman import HuffmanEncoder, ordered_counts
    rng = np.random.default_rng(11)
    for hi, n in ((2, 50), (37, 5000), (512, 20000), (1, 9)):
        v = rng.integers(0, hi, n)
        assert ordered_counts(v) == list(Counter(v.tolist()).items())
        assert ordered_counts(v.astype(np.int64) - 3) == list(Counter((v - 3).tolist()).items())   # negative symbols: generic path
        enc = HuffmanEncoder()
        enc.prepare_for_encoding(v)
        codes, widths = enc.encode_column(v)
        want = [enc.encode(int(x)) for x in v]
        assert codes.tolist() == [c for c, _ in want] and widths.tolist() == [w for _, w in want]
        codes, widths = IdentityEncoder(9).encode_column(v)
        assert codes.tolist() == v.tolist() and set(widths.tolist()) == {9}
    enc = HuffmanEncoder()
    enc.prepare_for_encoding(np.array([1, 2, 2, 5]))
    with pytest.raises(KeyError):
        enc.encode_column(np.array([1, 3]))


def test_bitbuffer_write_many_equals_single_writes():
    """The vectorised field writer the serialisers use leaves the same bytes and the same pending state as one
    `write` per field, whatever is already pending (widths 0..32, values wider than their field are masked)."""
    import random
    from bitbuffer import BitBuffer
    rnd = random.Random(7)
    for _ in range(60):
        fields = [(rnd.getrandbits(40), rnd.choice([0, 1, 1, 2, 3, 5, 7, 9, 16, 24, 31, 32])) for _ in range(rnd.randint(0, 400))]
        pre = [(rnd.getrandbits(12), rnd.randint(1, 12)) for _ in range(rnd.randint(0, 3))]
        a, b = io.BytesIO(), io.BytesIO()
        A, B = BitBuffer(a), BitBuffer(b)
        for v, w in pre:
            A.write(v, w)
            B.write(v, w)
        for v, w in fields:
            A.write(v, w)
        B.write_many([v for v, _ in fields], [w for _, w in fields])
        assert a.getvalue() == b.getvalue() and (A.buff, A.accumulated_size) == (B.buff, B.accumulated_size)
        A.flush()
        B.flush()
        assert a.getvalue() == b.getvalue()


def test_native_reader_rejects_hostile_headers():
    """Field widths and Huffman tables come from the file: widths beyond the bit buffer's 32-bit words, roots and child
    links outside the table or pointing backwards (a cycle) are refused instead of read."""
    import fic_b200
    data = bytes(64)
    ok_w, no_h = [8, 8, 5, 7], [-1, -1, -1, -1]
    empty = np.zeros((0, 3), np.int32)
    with pytest.raises(ValueError, match="bits wide"):
        fic_b200.tree_read_nodes(fic_b200.TREE_QUADTREE, data, 0, [3], [8, 8, 40, 7], no_h, empty, 64)
    with pytest.raises(ValueError, match="bits wide"):
        fic_b200.tree_read_nodes(fic_b200.TREE_HV, data, 0, None, [8, -1, 5, 7], no_h, empty, 64)
    loop = np.array([[0, 1, 0], [-1, -1, 3]], np.int32)            # row 0 links to itself
    with pytest.raises(ValueError, match="links to rows"):
        fic_b200.tree_read_nodes(fic_b200.TREE_QUADTREE, data, 0, [3], ok_w, [0, -1, -1, -1], loop, 64)
    far = np.array([[1, 7, 0], [-1, -1, 3]], np.int32)             # child outside the table
    with pytest.raises(ValueError, match="links to rows"):
        fic_b200.tree_read_nodes(fic_b200.TREE_QUADTREE, data, 0, [3], ok_w, [0, -1, -1, -1], far, 64)
    good = np.array([[1, 2, 0], [-1, -1, 3], [-1, -1, 4]], np.int32)
    with pytest.raises(ValueError, match="outside the table"):
        fic_b200.tree_read_nodes(fic_b200.TREE_QUADTREE, data, 0, [3], ok_w, [5, -1, -1, -1], good, 64)
