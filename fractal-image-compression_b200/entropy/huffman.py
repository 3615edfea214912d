"""Huffman coder of one coding dimension (reference: entropy/huffman.py:20-122).

Byte compatibility with the reference's files hinges on the tree SHAPE, which depends on how ties between equal
frequencies are broken: the reference pushes `Counter(symbols).items()` (first-occurrence order) through the standard
library `heapq` with nodes that compare by frequency only.  The same data structure with the same comparison is
used here, so the sift order -- and with it every code word -- is identical.
"""
from __future__ import annotations

import heapq
from collections import Counter, deque
from typing import Optional

import numpy as np

from bitbuffer import BitBuffer
from entropy.abccodec import Code, CodeWidth, EntropyEncoder

LEFT_BIT = 0
RIGHT_BIT = 1
INTERNAL_NODE_BIT = 0
LEAF_BIT = 1
UNAVAILABLE = -1
TABLE_VAL_BITS_ENCODING_SIZE_BITS = 8

CodingTable = dict[int, tuple[Code, CodeWidth]]


class HuffmanNode:
    """Tree node ordered by `freq` alone (the reference's dataclass(order=True) with compare=False on the rest)."""

    __slots__ = ("freq", "val", "left", "right")

    def __init__(self, freq: int, val: Optional[int], left: Optional["HuffmanNode"], right: Optional["HuffmanNode"]):
        self.freq = freq
        self.val = val
        self.left = left
        self.right = right

    def __lt__(self, other: "HuffmanNode") -> bool:
        return self.freq < other.freq

    def __le__(self, other: "HuffmanNode") -> bool:
        return self.freq <= other.freq

    def __eq__(self, other) -> bool:
        return isinstance(other, HuffmanNode) and self.freq == other.freq

    __hash__ = None

    @staticmethod
    def Leaf(val: int, freq: int) -> "HuffmanNode":
        return HuffmanNode(freq, val, None, None)

    @staticmethod
    def Inner(left: "HuffmanNode", right: "HuffmanNode") -> "HuffmanNode":
        return HuffmanNode(left.freq + right.freq, None, left, right)

    def is_leaf(self) -> bool:
        return self.left is None


def ordered_counts(symbols) -> list[tuple[int, int]]:
    """(symbol, frequency) in order of first occurrence -- what `Counter(symbols).items()` yields -- for a list or, without
    a Python-level pass over the elements, for a numpy array."""
    if isinstance(symbols, np.ndarray):
        n = symbols.size
        if n and symbols.dtype.kind in "iu" and int(symbols.min()) >= 0 and int(symbols.max()) < (1 << 20):
            # small non-negative symbols: histogram + first position by a reversed scatter (the last write wins), no sort
            # over the elements
            counts = np.bincount(symbols)
            first = np.full(counts.size, n, np.int64)
            first[symbols[::-1]] = np.arange(n - 1, -1, -1, dtype=np.int64)
            present = np.flatnonzero(counts)
            order = present[np.argsort(first[present], kind="stable")]
            return list(zip(order.tolist(), counts[order].tolist()))
        uniq, first, counts = np.unique(symbols, return_index=True, return_counts=True)
        order = np.argsort(first, kind="stable")
        return list(zip(uniq[order].tolist(), counts[order].tolist()))
    return list(Counter(symbols).items())


class HuffmanEncoder(EntropyEncoder):
    def __init__(self) -> None:
        self.root: Optional[HuffmanNode] = None
        self.coding_table: Optional[CodingTable] = None

    def prepare_for_encoding(self, symbols: list[int]) -> CodingTable:
        heap = [HuffmanNode.Leaf(sym, freq) for sym, freq in ordered_counts(symbols)]
        heapq.heapify(heap)
        while len(heap) > 1:
            first = heapq.heappop(heap)       # becomes the left child: code bit 0
            second = heapq.heappop(heap)
            heapq.heappush(heap, HuffmanNode.Inner(first, second))
        self.root = heap[0]
        self.coding_table = self._build_coding_table(self.root)
        return self.coding_table

    def encode(self, symbol: int) -> tuple[Code, CodeWidth]:
        return self.coding_table[symbol]

    def encode_column(self, values: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
        """Table lookup for a whole column (symbols are small non-negative integers: field values of a few bits)."""
        v = np.asarray(values)
        top = max(self.coding_table) if self.coding_table else -1
        if v.size == 0 or v.dtype.kind not in "iu" or top >= (1 << 20) or int(v.min()) < 0 or int(v.max()) > top:
            return super().encode_column(v)
        codes = np.zeros(top + 1, np.uint64)
        widths = np.full(top + 1, -1, np.int64)
        for sym, (code, width) in self.coding_table.items():
            codes[sym], widths[sym] = code, width
        w = widths[v]
        if (w < 0).any():
            raise KeyError(int(v[w < 0][0]))      # a symbol the tree was not built for, as `encode` would raise
        return codes[v], w

    def _build_coding_table(self, root: HuffmanNode) -> CodingTable:
        table: CodingTable = {}
        stack = [(root, 0, 0)]
        while stack:                          # a single-symbol alphabet gets the empty code (0, 0)
            node, code, depth = stack.pop()
            if node.is_leaf():
                table[node.val] = (code, depth)
            else:
                stack.append((node.right, (code << 1) | RIGHT_BIT, depth + 1))
                stack.append((node.left, (code << 1) | LEFT_BIT, depth + 1))
        return table


class HuffmanDecoder:
    """Walks the tree bit by bit (huffman.py:75-81).  A decoder made by the native file reader holds the tree as a
    table of (left, right, value) rows; the node objects are built when `root` (or `decode`) is first used."""

    def __init__(self, root: Optional[HuffmanNode] = None, table: Optional[np.ndarray] = None) -> None:
        self._root = root
        self._table = table

    @property
    def root(self) -> HuffmanNode:
        if self._root is None:
            rows = self._table.tolist()
            made = [HuffmanNode.Leaf(v, UNAVAILABLE) if left < 0 else HuffmanNode(UNAVAILABLE, None, None, None)
                    for left, _, v in rows]
            for node, (left, right, _) in zip(made, rows):
                if left >= 0:
                    node.left, node.right = made[left], made[right]
            self._root = made[0]
        return self._root

    @property
    def table(self) -> np.ndarray:
        """Rows (left, right, value), row 0 the root, left < 0 marking a leaf."""
        if self._table is None:
            index, order, stack = {}, [], [self._root]
            while stack:
                node = stack.pop()
                index[id(node)] = len(order)
                order.append(node)
                if node.left is not None:
                    stack.append(node.right)
                    stack.append(node.left)
            self._table = np.array([(-1, -1, int(n.val) if n.val is not None else 0) if n.left is None
                                    else (index[id(n.left)], index[id(n.right)], 0) for n in order], np.int32).reshape(-1, 3)
        return self._table

    def decode(self, read_buff: BitBuffer) -> int:
        node = self.root
        while not node.is_leaf():
            node = node.left if read_buff.read(1) == LEFT_BIT else node.right
        return node.val


class HuffmanSerializer:
    """Breadth-first dump: `0` for an inner node, `1` + value in `val_bits` bits for a leaf (huffman.py:83-98)."""

    def serialize(self, root: HuffmanNode, val_bits: int, write_buff: BitBuffer):
        pending = deque([root])
        while pending:
            node = pending.popleft()
            if node.is_leaf():
                write_buff.write(LEAF_BIT, 1)
                write_buff.write(node.val, val_bits)
            else:
                write_buff.write(INTERNAL_NODE_BIT, 1)
                pending.append(node.left)
                pending.append(node.right)


class HuffmanDeserializer:
    def deserialize(self, read_buff: BitBuffer, val_bits: int) -> HuffmanNode:
        def read_node() -> HuffmanNode:
            if read_buff.read(1) == LEAF_BIT:
                return HuffmanNode.Leaf(read_buff.read(val_bits), UNAVAILABLE)
            return HuffmanNode(UNAVAILABLE, None, None, None)

        root = read_node()
        if root.is_leaf() and root.val is not None:
            return root
        waiting = deque([root])              # inner nodes whose children are still to come, in stream order
        while waiting:
            parent = waiting.popleft()
            parent.left = read_node()
            parent.right = read_node()
            for child in (parent.left, parent.right):
                if child.val is None:
                    waiting.append(child)
        return root
