// `orfcompare` front-end (orfcompare.cpp:37-279 of the reference): loads a query and a template annotation
// that both carry CDSs, bundles them, and prints TX::score of every (query, template) pair of a bundle.
// The scoring itself (TX::score, Transcriptome.cpp:699-762) runs behind Backend::score_pairs; the host does
// the loading, the packing and the TSV.  With --reference the first / last codon of each CDS is printed too:
// six bases per transcript, read straight from the FASTA here.
#include "orfx_host.h"
#include "orfx_codon_table.h"
#include <algorithm>
#include <cstring>

namespace orfx {

namespace {

struct CmpDef { const char* name; bool flag; bool required; };
const CmpDef CMP_DEFS[] = {{"query", false, true}, {"template", false, true}, {"output", false, true}, {"threads", false, false},
                           {"use_id", true, false}, {"reference", false, false}, {"gpus", false, false}};

bool exists(const std::string& p) { FILE* f = fopen(p.c_str(), "r"); if (f) fclose(f); return f != nullptr; }

// codon_map[nts] of common.hpp:279 with the operator[] default for unknown keys
char codon_aa(const char* c3, size_t n) {
    static std::map<std::string, char> m;
    if (m.empty()) for (int i = 0; i < ORFX_CODON_TABLE_N; i++) m[std::string(ORFX_CODON_TABLE[i], 3)] = ORFX_CODON_TABLE[i][3];
    auto it = m.find(std::string(c3, n));
    return it == m.end() ? '\0' : it->second;
}
char complement(char c) { return nt_complement(c); }
// amino acid of the codon that starts at transcript offset `off` of `chain` (SEQ::translate on the spliced,
// upper-cased, strand-corrected sequence, common.hpp:477-495)
char aa_at(const Fasta& fa, const Fasta::Seq& chrom, const Chain& chain, char strand, int off) {
    char c3[3];
    const int L = chain_len(chain);
    size_t n = 0;
    for (int k = 0; k < 3 && off + k < L; k++) {
        const int g = nt2genome(chain, off + k, strand);
        char b = fa.at(chrom, g - 1);
        if (strand != '+') b = complement(b);
        c3[n++] = (char)toupper((unsigned char)b);
    }
    return codon_aa(c3, n);
}
// first and last amino acid of TX::load_seq()'s translation (Transcriptome.cpp:683-697)
void end_codons(const Fasta& fa, const Fasta::Seq& chrom, const Tx& t, char& first, char& last) {
    const int s5 = t.strand == '+' ? t.cds_start : t.cds_end, s3 = t.strand == '+' ? t.cds_end : t.cds_start;
    const int a = genome2nt(t.exons, s5, t.strand), b = genome2nt(t.exons, s3, t.strand);
    first = last = '\0';
    if (a < 0 || b < a + 2) return;
    const int naa = (b - a + 1) / 3;
    first = aa_at(fa, chrom, t.exons, t.strand, a);
    last = aa_at(fa, chrom, t.exons, t.strand, a + 3 * (naa - 1));
}

}  // namespace

bool parse_compare_cli(int argc, char** argv, Options& opt) {
    std::map<std::string, std::string> val;
    std::map<std::string, bool> set;
    for (int i = 1; i < argc; i++) {
        if (strlen(argv[i]) < 3 || argv[i][0] != '-' || argv[i][1] != '-') throw std::runtime_error(std::string("Unknown option ") + argv[i] + "!");
        const std::string name = argv[i] + 2;
        if (name == "help") {
            printf("orfcompare: run comparison of ORFs between two files.\n  --query GTF --template GTF --output TSV [--reference FASTA] [--use_id] [--threads N] [--gpus LIST]\n");
            return false;
        }
        const CmpDef* d = nullptr;
        for (const CmpDef& c : CMP_DEFS) if (name == c.name) d = &c;
        if (!d) throw std::runtime_error("Unknown option --" + name + "!");
        set[name] = true;
        if (!d->flag) {
            if (i + 1 >= argc) throw std::runtime_error("Option --" + name + " needs a value!");
            val[name] = argv[++i];
        }
    }
    for (const CmpDef& c : CMP_DEFS) if (c.required && !set.count(c.name)) throw std::runtime_error(std::string("Option --") + c.name + " required!");
    opt.query = val["query"];
    if (!exists(opt.query)) throw std::runtime_error("Query file does not exist! " + opt.query);
    opt.templates.assign(1, val["template"]);
    if (!exists(opt.templates[0])) throw std::runtime_error("Template file does not exist! " + opt.templates[0]);
    opt.use_id = set.count("use_id") > 0;
    if (set.count("reference")) {
        opt.reference = val["reference"];
        if (!exists(opt.reference)) throw std::runtime_error("Reference FASTA file does not exist!");
    }
    if (set.count("threads")) opt.threads = atoi(val["threads"].c_str());
    if (set.count("gpus")) { opt.devices.assign(1, atoi(val["gpus"].c_str())); }
    opt.output = val["output"];
    return true;
}

// replaces run() of orfcompare.cpp:37-171
int run_orfcompare(const Options& opt, const Backend& be) {
    if (!be.score_pairs) throw std::runtime_error("orfcompare needs a backend with pair scoring (orfx_score_pairs)");
    Fasta fasta;
    Transcriptome tr;
    if (!opt.reference.empty()) {
        fprintf(stderr, "loading reference genome\n");
        fasta.load(opt.reference);
        if (fasta.write_fai_if_missing(opt.reference)) fprintf(stderr, "FASTA index rebuilt.\n");
        tr.fasta = &fasta;
    }
    fprintf(stderr, "loading reference transcriptomes\n");
    tr.add(opt.templates[0], true, true);
    tr.add(opt.query, false, true);
    tr.build_cds_chains();
    fprintf(stderr, "sorting reference transcriptome\n");
    tr.set_cds_as_exons();
    tr.remove_non_coding();
    tr.sort(false);
    fprintf(stderr, "start removing duplicates\n");
    tr.deduplicate(opt.use_id);
    fprintf(stderr, "bundling transcriptome\n");
    tr.bundleup(opt.use_id, 0);
    fprintf(stderr, "starting main evaluation\n");

    FILE* out = fopen(opt.output.c_str(), "w");
    if (!out) throw std::runtime_error("Cannot open output file: " + opt.output);
    fputs("query_id\ttemplate_id\tquery_len\ttemplate_len\tlen_match\tlen_inframe\tlen_outframe\tlen_extra\tlen_missing\tlpi\tmlpi\tilpi\t"
          "query_start_codon\ttemplate_start_codon\tquery_stop_codon\ttemplate_stop_codon\n", out);

    // q->build_cds() / t->build_cds() of the loop (orfcompare.cpp:91,126), hoisted: both are idempotent
    for (Tx& t : tr.txs) t.build_cds();
    std::vector<int> active;
    for (int bi = 0; bi < (int)tr.bundles.size(); bi++) if (tr.bundles[bi].has_template && tr.bundles[bi].has_query) active.push_back(bi);
    TidClasses tid_classes;
    build_tid_classes(tr, tid_classes);
    PackedBatch pb;
    pack_bundles(tr, active, false, tid_classes, pb, be.host_alloc, be.host_free, false, PACK_COMPARE);
    orfx_params params;
    fill_params(opt, false, params);
    params.has_reference = 0;
    orfx_ctx* ctx = nullptr;
    if (be.create(&ctx, opt.devices.empty() ? 0 : opt.devices[0], &params)) { fclose(out); throw std::runtime_error(be.last_error()); }
    orfx_results res;
    const orfx_batch view = pb.view();
    const int rc = be.score_pairs(ctx, &view, &res);
    if (rc) { const std::string err = be.last_error(); be.destroy(ctx); fclose(out); throw std::runtime_error(err); }

    // first / last codons, once per transcript
    const bool ref = tr.fasta != nullptr;
    std::vector<char> c_first(tr.txs.size(), '-'), c_last(tr.txs.size(), '-');
    if (ref) {
        for (const Bundle& b : tr.bundles) {
            const Fasta::Seq* chrom = fasta.find(tr.seqnames[(size_t)b.seqid]);
            for (int i = 0; i < b.count; i++) end_codons(fasta, *chrom, tr.txs[(size_t)(b.first + i)], c_first[(size_t)(b.first + i)], c_last[(size_t)(b.first + i)]);
        }
    }
    int64_t row = 0;
    size_t next_active = 0;
    for (int bi = 0; bi < (int)tr.bundles.size(); bi++) {
        const Bundle& b = tr.bundles[(size_t)bi];
        const bool is_active = next_active < active.size() && active[next_active] == bi;
        if (is_active) next_active++;
        for (int i = 0; i < b.count; i++) {
            const size_t qg = (size_t)(b.first + i);
            const Tx& q = tr.txs[qg];
            if (q.is_templ) continue;
            if (!b.has_template) {
                fprintf(out, "%s\t-\t%d\t-\t-\t-\t-\t-\t-\t-\t-\t-\t", q.tid.c_str(), chain_len(q.cds));
                fputc(c_first[qg], out); fputs("\t-\t", out); fputc(c_last[qg], out); fputs("\t-\t\n", out);
                continue;
            }
            for (int j = 0; j < b.count; j++) {
                const size_t tg = (size_t)(b.first + j);
                const Tx& t = tr.txs[tg];
                if (!t.is_templ) continue;
                if (row >= res.n_rows || (size_t)pb.tx_global[(size_t)res.row_q_tx[row]] != qg || (size_t)pb.tx_global[(size_t)res.row_t_tx[row]] != tg) {
                    be.destroy(ctx); fclose(out);
                    throw std::runtime_error("orfcompare: pair order returned by the backend does not match the bundles");
                }
                fprintf(out, "%s\t%s\t%d\t%d\t%d\t%d\t%d\t%d\t%d\t%s\t%s\t%s\t", q.tid.c_str(), t.tid.c_str(), res.row_qlen[row], res.row_tlen[row],
                        res.row_match[row], res.row_inframe[row], res.row_outframe[row], res.row_extra[row], res.row_missing[row],
                        float_ostream(res.row_lpi[row]).c_str(), float_ostream(res.row_mlpi[row]).c_str(), float_ostream(res.row_ilpi[row]).c_str());
                fputc(c_first[qg], out); fputc('\t', out); fputc(c_first[tg], out); fputc('\t', out);
                fputc(c_last[qg], out); fputc('\t', out); fputc(c_last[tg], out); fputs("\t\n", out);
                row++;
            }
        }
    }
    be.destroy(ctx);
    fclose(out);
    fprintf(stderr, "Done\n");
    return 0;
}

}  // namespace orfx
