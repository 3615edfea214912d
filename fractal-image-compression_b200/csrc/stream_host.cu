// stream_host.cu -- host-side readers of the node section of the reference's container files, so that a file reaches the
// decoder kernels without one Python call per bit:
//   fic_tree_read_nodes   QuadtreeDeserializer._read_data (quadtree/serialization.py:182-223) and
//                         HVDeserializer._read_data (hvv2/serialization.py:186-224)
//   fic_huffman_read_tree HuffmanDeserializer.deserialize (entropy/huffman.py:100-124)
//   fic_pack_fields       BitBuffer.write applied to a whole column of fields (bitbuffer.py:18-37)
//   fic_hv_layout         the rectangle of every node of an HV preorder stream (hvv2/decoder.py:59-88 walks the same
//                         splits) and its depth -- what the decoder's leaf table and the writer's breadth-first order need
// No device work here; the file is a .cu only so that it builds with the rest of the library.
#include <deque>
#include <vector>

#include "fic_common.cuh"

namespace {

// MSB-first bit reader over the file's bytes (bitbuffer.py:39-56 reads big-endian 32-bit words, most significant bit
// first: the same bit order).  The reference can only read whole words, so the usable length is rounded down to 4 bytes.
struct BitReader {
    const uint8_t* data;
    int64_t nbits;
    int64_t pos;
    bool eof = false;
    uint32_t read(int width)
    {
        if (pos + width > nbits) { eof = true; return 0; }
        if (width == 0) return 0;
        if (width == 1) { const uint32_t b = (data[pos >> 3] >> (7 - (pos & 7))) & 1u; ++pos; return b; }
        const int64_t byte = pos >> 3;
        const int off = (int)(pos & 7);
        uint32_t v;
        if ((byte + 8) * 8 <= nbits) {   // eight bytes are there: one big-endian window, the field is cut out of it
            uint64_t w = 0;
            for (int k = 0; k < 8; ++k) w = (w << 8) | data[byte + k];
            v = (uint32_t)((w << off) >> (64 - width));
        } else {
            v = 0;
            for (int64_t p = pos; p < pos + width; ++p) v = (v << 1) | ((data[p >> 3] >> (7 - (p & 7))) & 1u);
        }
        pos += width;
        return v;
    }
};

struct DimCoder {
    int raw_width;          // used when root < 0
    int root;               // index into the flattened Huffman table, -1 = identity coding
    const int32_t* table;   // [m][3] = { left, right, value }; left < 0 marks a leaf
    uint32_t decode(BitReader& br) const
    {
        if (root < 0) return br.read(raw_width);
        int n = root;
        while (table[3 * n] >= 0 && !br.eof) n = br.read(1) == 0 ? table[3 * n] : table[3 * n + 1];   // huffman.py:75-81
        return (uint32_t)table[3 * n + 2];
    }
};

struct RawNode {
    int32_t kind, f[6];     // leaf: row, col, orientation, rotation, qs, qo;  HV inner: vertical, split
    int32_t child[4];
};

}  // namespace

extern "C" {

int fic_tree_read_nodes(int kind, const uint8_t* data, int64_t nbytes, int64_t bit_pos, const int32_t* root_exps, int n_roots,
                        const int32_t* raw_width, const int32_t* huff_root, const int32_t* huff_table, int huff_rows, int32_t* nodes,
                        int cap, int64_t* end_bit)
{
    FIC_REQUIRE(data && raw_width && huff_root && nodes && end_bit, "fic_tree_read_nodes: NULL argument");
    FIC_REQUIRE(kind == FIC_TREE_QUADTREE || kind == FIC_TREE_HV, "fic_tree_read_nodes: unknown tree kind %d", kind);
    FIC_REQUIRE(nbytes >= 0 && bit_pos >= 0 && n_roots >= 1 && (kind == FIC_TREE_HV || root_exps),
                "fic_tree_read_nodes: bad sizes");
    // untrusted input: field widths come from the file header (u8), table indices from a serialised tree
    FIC_REQUIRE(huff_rows >= 0 && (huff_rows == 0 || huff_table), "fic_tree_read_nodes: bad Huffman table");
    for (int r = 0; r < huff_rows; ++r) {   // children follow their parent in both layouts the library produces
        const int l = huff_table[3 * r], rt = huff_table[3 * r + 1];
        FIC_REQUIRE(l < 0 || (l > r && rt > r && l < huff_rows && rt < huff_rows),
                    "fic_tree_read_nodes: Huffman table row %d links to rows %d and %d of %d", r, l, rt, huff_rows);
    }
    for (int d = 0; d < 4; ++d) {
        FIC_REQUIRE(huff_root[d] >= 0 || (raw_width[d] >= 0 && raw_width[d] <= 32),
                    "fic_tree_read_nodes: field %d is %d bits wide; the bit buffer holds at most 32 (bitbuffer.py MAX_BIT_WIDTH)", d, raw_width[d]);
        FIC_REQUIRE(huff_root[d] < huff_rows, "fic_tree_read_nodes: field %d's Huffman root %d is outside the table of %d rows", d, huff_root[d], huff_rows);
    }
    const int arity = kind == FIC_TREE_QUADTREE ? 4 : 2;
    BitReader br{data, (nbytes / 4) * 32, bit_pos};
    DimCoder dim[4];
    for (int d = 0; d < 4; ++d) dim[d] = DimCoder{raw_width[d], huff_root[d], huff_table};
    int count = 0;
    std::vector<RawNode> tree;
    std::vector<int> stack;
    for (int r = 0; r < n_roots; ++r) {
        // one root: breadth-first stream; an inner node's children are the next unclaimed nodes, in order
        tree.clear();
        int64_t open_slots = 1;     // nodes still to be read for this root
        std::deque<int> waiting;    // inner nodes in stream order
        int filled = 0;             // children already attached to waiting.front()
        while (open_slots > 0) {
            RawNode nd{};
            const uint32_t leaf = br.read(1);
            if (leaf == 1) {
                nd.kind = 1;
                nd.f[0] = (int32_t)dim[0].decode(br);
                nd.f[1] = (int32_t)dim[1].decode(br);
                nd.f[2] = (int32_t)br.read(1);
                nd.f[3] = kind == FIC_TREE_QUADTREE ? (int32_t)br.read(2) : (int32_t)br.read(1) * 2;
                nd.f[4] = (int32_t)dim[2].decode(br);
                nd.f[5] = (int32_t)dim[3].decode(br);
            } else {
                nd.kind = 0;
                if (kind == FIC_TREE_HV) {
                    nd.f[0] = (int32_t)br.read(1);
                    nd.f[1] = (int32_t)br.read(16);
                }
            }
            FIC_REQUIRE(!br.eof, "Invalid size encoding, buffer returned EOF");
            const int id = (int)tree.size();
            tree.push_back(nd);
            --open_slots;
            if (id > 0) {
                tree[waiting.front()].child[filled++] = id;
                if (filled == arity) { waiting.pop_front(); filled = 0; }
            }
            if (nd.kind == 0) { waiting.push_back(id); open_slots += arity; }
        }
        // emit the root in preorder, the layout of fic_qt_encode / fic_hv_encode
        stack.clear();
        stack.push_back(0);
        std::vector<int> exps(tree.size(), 0);
        if (kind == FIC_TREE_QUADTREE) exps[0] = root_exps[r];
        while (!stack.empty()) {
            const int id = stack.back();
            stack.pop_back();
            const RawNode& nd = tree[id];
            if (count >= cap) { fic_set_error("fic_tree_read_nodes: node capacity %d too small", cap); return FIC_ECAP; }
            int32_t* o = nodes + 8 * (size_t)count++;
            o[0] = nd.kind;
            if (nd.kind == 1) {
                for (int k = 0; k < 6; ++k) o[1 + k] = nd.f[k];
            } else {
                o[1] = nd.f[0]; o[2] = nd.f[1]; o[3] = o[4] = o[5] = o[6] = 0;
            }
            o[7] = kind == FIC_TREE_QUADTREE ? exps[id] : 0;
            if (nd.kind == 0)
                for (int c = arity - 1; c >= 0; --c) {
                    exps[nd.child[c]] = exps[id] - 1;
                    stack.push_back(nd.child[c]);
                }
        }
    }
    *end_bit = br.pos;
    return count;
}

int fic_huffman_read_tree(const uint8_t* data, int64_t nbytes, int64_t bit_pos, int val_bits, int32_t* table, int cap,
                          int64_t* end_bit)
{
    FIC_REQUIRE(data && table && end_bit && nbytes >= 0 && bit_pos >= 0 && val_bits >= 0 && val_bits <= 32 && cap >= 1,
                "fic_huffman_read_tree: bad arguments");
    BitReader br{data, (nbytes / 4) * 32, bit_pos};
    // entropy/huffman.py:100-124: breadth-first, `1` + value for a leaf, `0` for an inner node whose two children follow
    // once every node read before them has had its turn
    auto read_node = [&](int k) {
        int32_t* row = table + 3 * (size_t)k;
        if (br.read(1) == 1) { row[0] = -1; row[1] = -1; row[2] = (int32_t)br.read(val_bits); }
        else { row[0] = 0; row[1] = 0; row[2] = 0; }
    };
    int count = 1;
    read_node(0);
    for (int k = 0; k < count && !br.eof; ++k) {
        if (table[3 * (size_t)k] < 0) continue;      // a leaf has no children to wait for
        if (count + 2 > cap) { fic_set_error("fic_huffman_read_tree: more than %d nodes", cap); return FIC_ECAP; }
        table[3 * (size_t)k] = count;
        table[3 * (size_t)k + 1] = count + 1;
        read_node(count);
        read_node(count + 1);
        count += 2;
    }
    FIC_REQUIRE(!br.eof, "Invalid size encoding, buffer returned EOF");
    *end_bit = br.pos;
    return count;
}

int64_t fic_pack_fields(const uint64_t* values, const int64_t* widths, int64_t n, uint32_t pending, int pending_bits,
                        uint8_t* out, int64_t out_cap, uint32_t* rest, int* rest_bits)
{
    if (!values || !widths || !out || !rest || !rest_bits || n < 0 || pending_bits < 0 || pending_bits > 31) {
        fic_set_error("fic_pack_fields: bad arguments");
        return FIC_EINVAL;
    }
    // bitbuffer.py:18-37: fields are appended most significant bit first; a big-endian 32-bit word leaves the moment it fills
    uint64_t acc = pending_bits ? (pending & ((1u << pending_bits) - 1u)) : 0;   // pending bits, right-aligned
    int have = pending_bits;
    int64_t written = 0;
    for (int64_t k = 0; k < n; ++k) {
        int w = (int)widths[k];
        if (w < 0 || w > 64) { fic_set_error("fic_pack_fields: field %lld has width %d", (long long)k, w); return FIC_EINVAL; }
        uint64_t v = values[k];
        while (w > 0) {
            const int take = w < 32 - have ? w : 32 - have;
            const uint64_t piece = (v >> (w - take)) & ((take == 64) ? ~0ull : ((1ull << take) - 1ull));
            acc = (acc << take) | piece;
            have += take;
            w -= take;
            if (have == 32) {
                if (written + 4 > out_cap) { fic_set_error("fic_pack_fields: output capacity too small"); return FIC_ECAP; }
                out[written++] = (uint8_t)(acc >> 24);
                out[written++] = (uint8_t)(acc >> 16);
                out[written++] = (uint8_t)(acc >> 8);
                out[written++] = (uint8_t)acc;
                acc = 0;
                have = 0;
            }
        }
    }
    *rest = (uint32_t)acc;
    *rest_bits = have;
    return written;
}

int fic_hv_layout(const int32_t* nodes, int n, int H, int W, int32_t* rects, int32_t* depth)
{
    FIC_REQUIRE(nodes && rects && depth && n >= 1 && H > 0 && W > 0, "fic_hv_layout: bad arguments");
    struct Open { int i, j, w, h, d; };
    std::vector<Open> stack;
    stack.push_back(Open{0, 0, W, H, 0});
    for (int k = 0; k < n; ++k) {
        FIC_REQUIRE(!stack.empty(), "fic_hv_layout: more nodes than the tree has room for (node %d)", k);
        const Open cur = stack.back();
        stack.pop_back();
        int32_t* r = rects + 4 * (size_t)k;
        r[0] = cur.i; r[1] = cur.j; r[2] = cur.w; r[3] = cur.h;
        depth[k] = cur.d;
        const int32_t* nd = nodes + 8 * (size_t)k;
        if (nd[0] == 0) {
            const int s = nd[2] + 1;   // split_idx is the last column / row of the first child
            if (nd[1]) {   // vertical split: left | right (hvv2/decoder.py:83-85)
                FIC_REQUIRE(s > 0 && s < cur.w, "fic_hv_layout: node %d splits a width of %d after column %d", k, cur.w, s - 1);
                stack.push_back(Open{cur.i, cur.j + s, cur.w - s, cur.h, cur.d + 1});
                stack.push_back(Open{cur.i, cur.j, s, cur.h, cur.d + 1});
            } else {       // horizontal split: top / bottom (hvv2/decoder.py:86-88)
                FIC_REQUIRE(s > 0 && s < cur.h, "fic_hv_layout: node %d splits a height of %d after row %d", k, cur.h, s - 1);
                stack.push_back(Open{cur.i + s, cur.j, cur.w, cur.h - s, cur.d + 1});
                stack.push_back(Open{cur.i, cur.j, cur.w, s, cur.d + 1});
            }
        }
    }
    FIC_REQUIRE(stack.empty(), "fic_hv_layout: the stream ends before the tree is complete");
    return FIC_OK;
}

}  // extern "C"
