"""Generates the binarisation-tree tables of the macroblock-level syntax elements (avr_walk.cuh, Walker::tree).

A node is one bin: its context (as an offset from the element's base context, one value for P / B slices and one for I
slices where they differ; 0xFF = the bin is end_of_slice-style `terminate`) and where to go on 0 / on 1 -- another
node, or a leaf (0x80 | value).  Node 0 is the root.  Word = ctx_other | ctx_islice << 8 | next0 << 16 | next1 << 24.

    python tools/gen_trees.py        # prints the C arrays pasted into avr_walk.cuh

Trees follow libavcodec's decoders: decode_cabac_intra_mb_type h264_cabac.c:1300-1330, P mb_type :2008-2023,
B mb_type :1973-2007, decode_cabac_p_mb_sub_type :1445-1454, decode_cabac_b_mb_sub_type :1455-1471,
decode_cabac_mb_chroma_pre_mode :1383-1406."""
TERM = 0xFF


class Node:
    def __init__(self, ctx, zero, one, ctx_i=None):
        self.ctx, self.ctx_i, self.zero, self.one = ctx, ctx if ctx_i is None else ctx_i, zero, one


def leaf(v):
    assert 0 <= v < 0x80
    return ("leaf", v)


def flatten(root):
    nodes, order = [], []

    def visit(n):
        n.index = len(order)
        order.append(n)
        for b in (n.zero, n.one):
            if isinstance(b, Node):
                visit(b)
    visit(root)
    for n in order:
        nx = [(0x80 | b[1]) if isinstance(b, tuple) else b.index for b in (n.zero, n.one)]
        assert all(x < 0x100 for x in nx) and all(not isinstance(b, Node) or b.index < 0x80 for b in (n.zero, n.one))
        nodes.append(n.ctx | n.ctx_i << 8 | nx[0] << 16 | nx[1] << 24)
    return nodes


def bits_tree(ctxs, value_of, prefix=()):
    """A complete tree over len(ctxs) bins (most significant first); value_of(bits tuple) -> leaf value or a Node."""
    if len(prefix) == len(ctxs):
        v = value_of(prefix)
        return v if isinstance(v, Node) else leaf(v)
    c = ctxs[len(prefix)]
    return Node(c[0], bits_tree(ctxs, value_of, prefix + (0,)), bits_tree(ctxs, value_of, prefix + (1,)), c[1])


INTRA_LEAF = 100   # "this is an intra macroblock: read the intra mb_type next"


def p_mb_type():       # base 14; leaves: ptype 0..3 (shape = ptype + 1), INTRA_LEAF
    return Node(0, Node(1, Node(2, leaf(0), leaf(3)), Node(3, leaf(2), leaf(1))), leaf(INTRA_LEAF))


def b_mb_type():       # base 27, root context chosen by the caller; leaves t = 0..22, INTRA_LEAF
    def four(bits):    # bits = bin(27+4) bin(27+5) bin(27+5) bin(27+5)
        v = bits[0] << 3 | bits[1] << 2 | bits[2] << 1 | bits[3]
        if v < 8:
            return v + 3
        if v == 13:
            return INTRA_LEAF
        if v == 14:
            return 11
        if v == 15:
            return 22
        return Node(5, leaf(2 * v - 4), leaf(2 * v + 1 - 4))
    deep = bits_tree([(4, 4), (5, 5), (5, 5), (5, 5)], four)
    return Node(0, leaf(0), Node(3, Node(5, leaf(1), leaf(2)), deep))


def intra_mb_type():   # contexts relative to base (+2 in I slices, done by the caller); `is` = 1 in I slices
    def rest(b_a):     # after bin A
        def tail(add):  # bins D (x2) and E (x1)
            return Node(3, Node(3, leaf(1 + 12 * b_a + add), leaf(2 + 12 * b_a + add), 5),
                        Node(3, leaf(3 + 12 * b_a + add), leaf(4 + 12 * b_a + add), 5), 4)
        return Node(2, tail(0), Node(2, tail(4), tail(8), 3))
    # root: context chosen by the caller; then terminate (1 = I_PCM, leaf 25); then bin A at base + 1
    return Node(0, leaf(0), Node(TERM, Node(1, rest(0), rest(1)), leaf(25)))


def p_sub_type():      # base 21; leaves sub = 0..3
    return Node(0, Node(1, leaf(1), Node(2, leaf(3), leaf(2))), leaf(0))


def b_sub_type():      # base 36; leaves t = 0..12
    def two(base_t):   # t = base_t + 2 * bin(39) + bin(39)
        return Node(3, Node(3, leaf(base_t), leaf(base_t + 1)), Node(3, leaf(base_t + 2), leaf(base_t + 3)))
    return Node(0, leaf(0), Node(1, Node(3, leaf(1), leaf(2)),
                                 Node(2, two(3), Node(3, two(7), Node(3, leaf(11), leaf(12))))))


def chroma_pred():     # base 64, root context chosen by the caller; leaves 0..3
    return Node(0, leaf(0), Node(3, leaf(1), Node(3, leaf(2), leaf(3))))


def one_bin():         # a single decision whose context the caller chooses
    return Node(0, leaf(0), leaf(1))


def one_terminate():
    return Node(TERM, leaf(0), leaf(1))


def main():
    trees = [("k_tree_p_mb", p_mb_type()), ("k_tree_b_mb", b_mb_type()), ("k_tree_intra_mb", intra_mb_type()),
             ("k_tree_p_sub", p_sub_type()), ("k_tree_b_sub", b_sub_type()), ("k_tree_chroma_pred", chroma_pred()),
             ("k_tree_one", one_bin()), ("k_tree_term", one_terminate())]
    for name, t in trees:
        w = flatten(t)
        print("AVR_TABLE_QUAL uint32_t %s[%d] = {%s};" % (name, len(w), ", ".join("0x%08Xu" % x for x in w)))


if __name__ == "__main__":
    main()
