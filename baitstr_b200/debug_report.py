"""``-d/--debug``: the per-read stderr report of the reference's ProcessSequence
(/root/reference/src/select_STR_reads.pyx:294-311) with the messages of its two matchers
(:76-87 default algorithm, :138-140 and :157-163 ``-x``).

The regex hits and their verdicts are enumerated on the GPU (``bstr_debug_reads``,
csrc/matcher.cuh: debug_read); this module only turns the codes into the reference's lines.
stdout is never affected.  With more than one worker process the reference's blocks interleave
nondeterministically; the order here is the input order (= the reference with ``-p 1``).
"""
RULE = "-" * 79                                                                   # pyx:297
_LINES = {ord("N"): "N in motif.", ord("5"): "Insufficient 5' flank.", ord("3"): "Insufficient 3' flank."}


def render_read(name, info, fwd_codes, rc_codes):
    """One read's block.  info = 8 integers as written by bstr_debug_reads."""
    out = [RULE, name]
    for strand, codes in ((0, fwd_codes), (1, rc_codes)):
        count, is_none, tail_code, tail_value = (int(x) for x in info[4 * strand:4 * strand + 4])
        out.extend(_LINES[c] for c in bytes(codes[:count]))
        if tail_code:                                                             # pyx:157-163
            out.append("Insufficient %s' flank.: %d" % (chr(tail_code), tail_value))
        if is_none:
            out.append("0 valid matches.")                                        # pyx:303-304, 310-311
            break
    return "\n".join(out) + "\n"
