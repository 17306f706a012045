"""Stats-TSV rendering of a Results object, for comparing C-ABI level results with the reference's
--stats output without going through the CLI (mirrors orfanage.cpp:793-814 / common.hpp:101-120)."""
import numpy as np

from orfanage_b200 import STATUS_NOTES

HEADER = ("query_id\ttemplate_id\tsegment\tnotes\tnum_templates\tquery_len\ttemplate_len\tunion_len\tpass\tlen_match\t"
          "len_inframe\tlen_outframe\tlen_extra\tlen_missing\tlength_pi\tmatch_length_pi\tinframe_length_pi\t"
          "alignment_match\tstart_match\tstop_match\tpi")
KEEP19 = "-\t-\tkeep_cds\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-\t-"


def g6(v):
    """operator<<(ostream&, float) at default precision == printf("%g")."""
    return "%g" % float(np.float32(v))


def stats_lines(res, tx_id, mode_all=False):
    """Lines (without header) the reference prints for the queries of an active bundle set."""
    out = []
    r = res.rows
    for b in res.best:
        qname = tx_id[b["q_tx"]]
        if b["t_tx"] < 0:
            out.append(qname + "\t" + KEEP19)
            continue
        o0, o1 = int(b["row_off"]), int(b["row_off"]) + int(b["n_rows"])
        o = o0
        while o < o1:
            front = int(r["print_order"][o])
            ks, ke = r["cds_start"][front], r["cds_end"][front]
            oe = o
            while oe < o1 and r["cds_start"][r["print_order"][oe]] == ks and r["cds_end"][r["print_order"][oe]] == ke:
                oe += 1
            tid = qname + (".%d-%d" % (ks, ke) if mode_all else "")
            front_note = None
            if mode_all:
                if r["pass"][front]:
                    front_note = "best_gtf" if front == b["best_row"] else "gtf"
            elif front == b["best_row"]:
                front_note = "best_gtf"
            for k in range(o, oe):
                row = int(r["print_order"][k])
                note = front_note if (k == o and front_note) else STATUS_NOTES[r["status"][row]]
                if r["status"][row] == 1:
                    seg = "4294967295-4294967295"
                else:
                    seg = "%d-%d" % (np.uint32(r["seg_start"][row]), np.uint32(r["seg_end"][row]))
                f = [tid, tx_id[r["t_tx"][row]], seg, note, str(int(r["n_dup"][row]) + 1), str(r["qlen"][row]), str(r["tlen"][row]),
                     str(r["ulen"][row]), str(int(r["pass"][row])), str(r["match"][row]), str(r["inframe"][row]),
                     str(r["outframe"][row]), str(r["extra"][row]), str(r["missing"][row]), g6(r["lpi"][row]), g6(r["mlpi"][row]),
                     g6(r["ilpi"][row]), "0", str(int(r["start_match"][row])), "0", "0"]
                out.append("\t".join(f))
            o = oe
    return out


def compare_lines(r, names):
    """orfcompare's first 12 TSV columns for the pairs returned by orfx_score_pairs (floats as operator<< prints them)."""
    out = []
    for i in range(len(r["q_tx"])):
        out.append([names[r["q_tx"][i]], names[r["t_tx"][i]]] + [str(int(r[k][i])) for k in ("qlen", "tlen", "match", "inframe", "outframe",
                   "extra", "missing")] + ["%g" % r[k][i] for k in ("lpi", "mlpi", "ilpi")])
    return out
