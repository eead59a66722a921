"""``-c/--check FILE``: the reference's stderr report on a list of expected STRs.

Host-side restatement of /root/reference/src/select_STR_reads.pyx:12-49 (``MotifDetails``,
``ReadSTRList``), :191-208 (``PercentIdentity``), :209-286 (``RegisterSTR``) and :287-292
(``PrintPerformanceLog``).  It never changes stdout.  Its input is the annotated FASTQ the GPU
pipeline produced (the header carries both strands' motif, copies, start and end, the next
line the upper-cased read), so no matching happens here.

What the reference prints, and this module reproduces (checked against the live reference,
tests/golden/check_strs.json):

* per selected read whose flanks agree with no listed STR by more than 90 %:
  ``No match for <motif of the reverse strand>:<copies of the forward strand>:<lflank>:<rflank>.
  Best pid:<..>`` -- the mix of strands is the reference's (it reuses ``motif`` at pyx:250 and
  keeps ``match[1]`` at pyx:285);
* after the last file, one ``Found 0 reads supporting <motif>:<copies>:<lflank>:<rflank>`` per
  listed STR, in file order: the reference increments ``found`` in its pool workers' copies of
  the list (pyx:356-357), so its parent always reports 0.  ``found`` is counted here as well
  (``Entry.found``) but the printed figure follows the reference;
* the left flank is ``string[start-20:start]`` with a NEGATIVE lower bound when the STR starts
  before base 20 (pyx:217 compares instead of assigning), i.e. Python slices from the end;
* both flanks empty raises ZeroDivisionError (pyx:207), as in the reference.
"""
import sys
from itertools import zip_longest

_COMPLEMENT = str.maketrans("atcgnATCGN", "tagcnTAGCN")                       # pyx:188


def _revcomp(seq):
    return seq.translate(_COMPLEMENT)[::-1]


class Entry(object):
    """One line of the check file (pyx:12-15, :42-47)."""

    def __init__(self, lflank, rflank):
        self.lflank, self.rflank, self.found = lflank, rflank, 0


def read_str_list(filename):
    """pyx:17-49: ``motif copies lflank rflank`` per line -> {(motif, copies): [Entry, ...]}."""
    strlist = {}
    if filename is None:
        return strlist
    with open(filename, "r") as fh:
        for line in fh:
            motif, copies, lflank, rflank = line.strip().split()
            strlist.setdefault((motif, int(copies)), []).append(Entry(lflank, rflank))
    return strlist


def percent_identity(entry, lflank, rflank):
    """pyx:191-207: left flanks compared from the STR outwards, right flanks from the STR
    onwards, over the shorter of the two, case-insensitive."""
    same = diff = 0
    for r, q in zip_longest(entry.lflank[::-1], lflank[::-1]):
        if r is not None and q is not None:
            if r.upper() == q.upper():
                same += 1
            else:
                diff += 1
    for r, q in zip_longest(entry.rflank, rflank):
        if r is not None and q is not None:
            if r.upper() == q.upper():
                same += 1
            else:
                diff += 1
    return (same * 100.0) / (same + diff)


def _flanks(string, start, end):
    lo = start - 20                         # pyx:216-217: stays negative (`s1 == 0` is a comparison)
    hi = min(end + 20, len(string))         # pyx:219-221
    return string[lo:start], string[end:hi]


def register_str(string, match, rc_match, to_check, err=None):
    """pyx:209-286.  match / rc_match: (motif, copies, start, end), rc_match in the coordinates of
    the reverse-complemented read.  Returns True when a listed STR was found."""
    err = sys.stderr if err is None else err
    max_pid = 0
    lflank = rflank = motif = ""
    # forward strand first (pyx:211-244), then the reverse complement (pyx:246-283)
    for strand, (motif, copies, start, end) in ((string, match), (_revcomp(string), rc_match)):
        lflank, rflank = _flanks(strand, start, end)
        for key, lf, rf in (((motif, copies), lflank, rflank),
                            ((_revcomp(motif), copies), _revcomp(rflank), _revcomp(lflank))):
            for entry in to_check.get(key, ()):
                pid = percent_identity(entry, lf, rf)
                if pid > 90:
                    entry.found += 1
                    return True
                if pid > max_pid:
                    max_pid = pid
    err.write("No match for %s:%d:%s:%s. Best pid:%2.2f\n" % (motif, match[1], lflank, rflank, max_pid))
    return False


def annotated_fastq_lines(data, to_check):
    """register_str over every record of an annotated FASTQ buffer (bytes) as written by the
    select stage (``@name\\tfmotif\\tfcopies\\tfstart\\tfend\\trmotif\\trcopies\\trstart\\trend``):
    yields, per record, the "No match" line or "" when a listed STR was found."""
    import io
    lines = data.split(b"\n")
    for i in range(0, len(lines) - 3, 4):
        fields = lines[i].decode("latin-1").rsplit("\t", 8)       # the name itself may hold tabs
        seq = lines[i + 1].decode("latin-1")
        match = (fields[1], int(fields[2]), int(fields[3]), int(fields[4]))
        rc_match = (fields[5], int(fields[6]), int(fields[7]), int(fields[8]))
        buf = io.StringIO()
        register_str(seq, match, rc_match, to_check, buf)
        yield buf.getvalue()


def register_annotated_fastq(data, to_check, err=None):
    """All "No match" lines of an annotated FASTQ buffer, written to `err`."""
    err = sys.stderr if err is None else err
    for line in annotated_fastq_lines(data, to_check):
        err.write(line)


def print_performance_log(to_check, err=None):
    """pyx:287-292, with the reference's observable count (always 0, see the module docstring)."""
    err = sys.stderr if err is None else err
    for (motif, copies), entries in to_check.items():
        for entry in entries:
            err.write("Found %d reads supporting %s:%d:%s:%s\n" % (0, motif, copies, entry.lflank, entry.rflank))
