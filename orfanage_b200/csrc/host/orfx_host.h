// Host side of the B200-native ORFanage hot path: the reference's surface (CLI flags, GTF/FASTA
// loading, template preparation, locus bundling, GTF / stats emission) around the C-ABI in
// include/orfx.h.  Nothing here scores or rescues an ORF: that is the device library's job.
#pragma once
#include <cstdint>
#include <cstdio>
#include <map>
#include <string>
#include <vector>
#include <stdexcept>
#include "orfx.h"

namespace orfx {

// --- interval chains (host mirror of SEGTP/CHAIN, Transcriptome.h:23-563) -------------------
struct Seg {
    int32_t s = 0, e = 0;
    int8_t phase = 0;
    int32_t len() const { return e + 1 - s; }
};
typedef std::vector<Seg> Chain;

int chain_len(const Chain& c);
// transcript-order offset (5'->3') of genomic position g, -1 when g is intronic/outside
// (CHAIN::genome2nt, Transcriptome.h:447-462)
int genome2nt(const Chain& c, int g, char strand);
// genomic position of transcript-order offset; clamps to the 3' end (CHAIN::nt2genome, h:463-481)
int nt2genome(const Chain& c, int off, char strand);
// GTF phase column for every segment, 5'->3' (CHAIN::assign_phase, h:432-446)
void assign_phase(Chain& c, char strand, int start_phase);
// keep the part of c inside [lo,hi] (CHAIN::cut, h:268-289)
void cut_chain(Chain& c, int lo, int hi);

// --- transcripts -----------------------------------------------------------------------------
struct Tx {
    bool is_templ = false;
    int id = -1;
    std::string tid, gid;
    int seqid = -1;
    char strand = '.';
    int cds_start = 0, cds_end = 0, cds_phase = 0;
    bool is_coding = false;
    std::string source, cds_source;
    std::map<std::string, std::string> attrs;
    Chain exons, cds;

    int start() const { return exons.front().s; }
    int end() const { return exons.back().e; }
    void build_cds();            // TX::build_cds, Transcriptome.cpp:92-122
    void correct_chain_len();    // TX::correct_chain_len, cpp:123-130
    void remove_cds();           // TX::remove_cds, cpp:131-138
    bool add_attribute(const std::string& k, const std::string& v) { return attrs.insert(std::make_pair(k, v)).second; }
    std::string gtf(const std::string& seqname) const;   // TX::str, Transcriptome.h:734-783
};
bool tx_less(const Tx& a, const Tx& b);    // TX::operator<, Transcriptome.h:668-684
bool tx_same(const Tx& a, const Tx& b);    // TX::operator==, Transcriptome.h:703-722

struct GtfRecord {   // one transcript as read from a GTF (what the reference gets from gclib's GffObj)
    std::string seqname, source, tid, gid;
    bool has_gid = false, has_cds = false;
    char strand = '.';
    std::vector<std::pair<uint32_t, uint32_t>> exons;   // ascending
    uint32_t cd_start = 0, cd_end = 0;
    char cd_phase = '.';
    std::vector<std::pair<std::string, std::string>> attrs;   // transcript-line attributes, file order
};
void read_gtf(const std::string& path, std::vector<GtfRecord>& out);
// Records converted to transcripts, not yet part of a Transcriptome: Tx::seqid indexes `seqnames` (the names of all records of
// the file, first-seen order), Tx::id counts from 0.
struct ParsedTxs { std::vector<Tx> txs; std::vector<std::string> seqnames; };
void convert_records(const std::vector<GtfRecord>& recs, bool is_templ, bool coding_only, ParsedTxs& out);

// --- FASTA ------------------------------------------------------------------------------------
// The FASTA is memory-mapped and read through its samtools-style index (the .fai next to it when present, else
// one scan that also writes it, as orfanage.cpp:989-1011 does): only the bases under the bundles are ever touched,
// a 3 Gb genome is not copied into memory.
class Fasta {
public:
    struct Seq {
        std::string name;
        int64_t length = 0, file_off = 0;
        int line_bases = 0, line_bytes = 0;
        std::string mem;          // records with ragged lines are held in memory instead
        bool in_mem = false;
    };
    Fasta() {}
    Fasta(const Fasta&) = delete;
    Fasta& operator=(const Fasta&) = delete;
    ~Fasta();
    void load(const std::string& path);
    const Seq* find(const std::string& name) const;
    // copies bases [pos0, pos0 + len) (0-based) of a record into dst
    void fetch(const Seq& s, int64_t pos0, int64_t len, char* dst) const;
    char at(const Seq& s, int64_t pos0) const { char c; fetch(s, pos0, 1, &c); return c; }
    bool write_fai_if_missing(const std::string& path) const;
private:
    bool read_fai(const std::string& fai);
    void scan();
    std::vector<Seq> recs_;
    std::map<std::string, size_t> idx_;
    const char* data_ = nullptr;
    size_t size_ = 0;
    bool fai_was_read_ = false;
};

struct Bundle {
    int first = 0, count = 0;        // range in Transcriptome::txs
    int seqid = -1;
    char strand = 0;
    int bstart = INT32_MAX, bend = 0;
    std::string gid;
    bool has_template = false, has_query = false;
};

class Transcriptome {
public:
    std::vector<Tx> txs;
    std::vector<Bundle> bundles;
    std::vector<std::string> seqnames;
    const Fasta* fasta = nullptr;

    void add(const std::string& gtf_path, bool is_templ, bool coding_only);   // Transcriptome::add, cpp:866-894
    void add_records(const std::vector<GtfRecord>& recs, bool is_templ, bool coding_only);   // the same from parsed records
    void append(ParsedTxs&& p);       // converted records join the transcriptome (sequence ids mapped, transcript ids assigned)
    void build_cds_chains();          // cpp:1033-1044
    void correct_chain_len();         // cpp:1045-1051
    void sort(bool use_id);           // Transcriptome.h:904-924
    void set_cds_as_exons();          // cpp:1141-1145
    void remove_non_coding();         // cpp:1146-1150
    size_t deduplicate(bool use_id);  // cpp:1124-1140
    size_t bundleup(bool use_id, uint32_t overhang);   // cpp:920-1032
    size_t clean_short_orfs(int minlen);   // cpp:1052-1061
private:
    std::map<std::string, int> seqids_;
};

// --- options (GlobalParameters, orfanage.cpp:99-142) -------------------------------------------
struct Options {
    bool clean_query = false, clean_templ = false, rescue = false, use_id = false, non_aug = false;
    bool keep_all_cds = false, keep_cds_if_not_found = false, spliced_overhang = false;
    int lpi = -1, ilpi = -1, mlpi = -1, minlen = -1, overhang = 0, threads = 1;
    int pi = 0;   // --pi: != 0 enables the protein alignment of every scored candidate (orfx_params::align_enabled)
    std::vector<int> modes{ORFX_MODE_LONGEST_MATCH, ORFX_MODE_BEST, ORFX_MODE_START_MATCH,
                           ORFX_MODE_LONGEST, ORFX_MODE_FIRST, ORFX_MODE_ALL};
    std::string reference, query, output, stats;
    std::vector<std::string> templates;
    bool stats_set = false;
    // build-specific
    std::vector<int> devices{0};
    bool timing = false;
};
// Parses argv with the reference's rules (arg_parse.hpp:101-209; orfanage.cpp:835-1098).
// Returns false when the program should exit 0 (help).
bool parse_cli(int argc, char** argv, Options& opt);
void fill_params(const Options& opt, bool want_rows, orfx_params& p);

// --- batcher: bundles -> flat arrays ------------------------------------------
struct PackedBatch {
    std::vector<int> bundle_ids;            // index into Transcriptome::bundles
    std::vector<int> tx_global;             // batch tx index -> Transcriptome::txs index
    // the flat arrays live in one arena obtained from `alloc` (pinned when the backend offers it)
    void* arena = nullptr;
    size_t arena_bytes = 0;
    void (*dealloc)(void*) = nullptr;
    uint8_t* seq = nullptr;
    int64_t* seq_off = nullptr;
    int32_t *bundle_start = nullptr, *bundle_tx_off = nullptr;
    uint8_t *tx_kind = nullptr, *tx_strand = nullptr;
    int32_t *tx_seg_off = nullptr, *tx_cds_start = nullptr, *tx_cds_end = nullptr, *tx_tid_class = nullptr;
    orfx_seg* segs = nullptr;
    int8_t* seg_phase = nullptr;
    int32_t* seg_seq_off = nullptr;         // islands layout only
    int32_t* tx_ext_lo = nullptr;           // --overhang only: usable flank below / above each query's exon chain
    int32_t* tx_ext_hi = nullptr;
    uint8_t* tx_ext_flags = nullptr;        // ORFX_EXT_* (a flank that re-reads the chromosome's edge base)
    // segment dictionary (orfx_batch::seg_ref): second arena from the same allocator; NULL when not built
    void* dict_arena = nullptr;
    uint16_t* seg_ref = nullptr;
    int32_t* bundle_dict_off = nullptr;
    orfx_seg* dict_segs = nullptr;
    int8_t* dict_phase = nullptr;
    int32_t* dict_seq_off = nullptr;
    int64_t n_dict = 0;
    std::vector<int64_t> seq_exc_pos;       // ORFX_SEQ_NT2: the bases that are none of A C G T
    std::vector<uint8_t> seq_exc_code;
    int32_t n_tx = 0;
    int64_t n_segs = 0, seq_bytes = 0, n_queries = 0;   // seq_bytes counts bases
    int32_t seq_format = ORFX_SEQ_RAW;
    PackedBatch() {}
    PackedBatch(const PackedBatch&) = delete;
    PackedBatch& operator=(const PackedBatch&) = delete;
    ~PackedBatch() { reset(); }
    void reset();
    orfx_batch view() const;
};
// Packs the listed bundles (each must hold a template and a query).  mode = PACK_CLEAN builds a
// *cleaning batch* instead (orfx_clean_templates in include/orfx.h): every coding transcript with a CDS
// chain is packed as ORFX_TX_QUERY with its exon chain and CDS bounds, everything else as ORFX_TX_SKIP.
// mode = PACK_COMPARE builds a *compare batch* (orfx_score_pairs): queries and templates both carry their
// CDS chain, no sequence.
// equal transcript_id strings <=> equal value, one entry per transcript of tr.txs (orfx_batch::tx_tid_class)
typedef std::vector<int> TidClasses;
enum { PACK_EVAL = 0, PACK_CLEAN = 1, PACK_COMPARE = 2 };
char nt_complement(char c);   // ntComplement (gclib), IUPAC, case preserving
void pack_bundles(const Transcriptome& tr, const std::vector<int>& bundle_ids, bool keep_all_cds,
                  const TidClasses& tid_classes, PackedBatch& out,
                  void* (*alloc)(size_t) = nullptr, void (*dealloc)(void*) = nullptr, bool islands = true, int mode = PACK_EVAL,
                  bool nt4 = false,    // nt4: islands as 4-bit base codes (ORFX_SEQ_NT4), half the upload
                  int overhang = 0, bool spliced_overhang = false,   // --overhang flanks (TX::extend_seq, Transcriptome.cpp:342-399)
                  bool template_seq = false,    // islands also cover the template CDS segments (orfx_params::align_enabled)
                  bool nt2 = false,             // islands as 2-bit base codes + exception list (ORFX_SEQ_NT2), a quarter of the raw upload
                  bool dict = false);           // also build the segment dictionary (orfx_batch::seg_ref): the upload then skips the per-segment arrays
// Per bundle, the distinct (segment, phase, sequence offset) records of its transcripts and a 16-bit reference per segment.
// Returns false (and leaves the batch without a dictionary) when a bundle has more than 65536 distinct records.
bool build_segment_dictionary(PackedBatch& out, void* (*alloc)(size_t));
void build_tid_classes(const Transcriptome& tr, TidClasses& out);
// greedy longest-processing-time split of active bundles over n shards
void shard_bundles(const Transcriptome& tr, const std::vector<int>& active, int n, std::vector<std::vector<int>>& shards);

// --- backend: the five C-ABI entry points the runner drives -----------------------------------
struct Backend {
    int (*create)(orfx_ctx**, int, const orfx_params*);
    int (*submit)(orfx_ctx*, const orfx_batch*);
    int (*wait)(orfx_ctx*, orfx_results*);
    int (*release)(orfx_ctx*);
    void (*destroy)(orfx_ctx*);
    const char* (*last_error)();
    void* (*host_alloc)(size_t);      // optional (may be NULL): pinned host memory for the packed arrays
    void (*host_free)(void*);
    // optional (may be NULL): template cleaning, --cleant / --rescue (orfx_clean_templates)
    int (*clean)(orfx_ctx*, const orfx_batch*, int, orfx_clean_out*, int64_t*);
    // optional (may be NULL): pair scoring for the orfcompare front-end (orfx_score_pairs)
    int (*score_pairs)(orfx_ctx*, const orfx_batch*, orfx_results*);
};

// load + template preparation + bundling (orfanage.cpp:313-367)
// `be` is needed only for --cleant (Transcriptome::clean_cds runs behind Backend::clean)
void prepare_transcriptome(const Options& opt, Fasta& fasta, Transcriptome& tr, bool verbose, const Backend* be = nullptr);
// replaces run() (orfanage.cpp:307-822): prepares, bundles, evaluates through `be`, writes outputs
int run_orfanage(const Options& opt, const Backend& be);
// the companion tool (orfcompare.cpp): argument parsing and run()
bool parse_compare_cli(int argc, char** argv, Options& opt);
int run_orfcompare(const Options& opt, const Backend& be);

// emission helpers (orfanage.cpp:662-815)
struct Emitter {
    FILE* gtf = nullptr;
    FILE* stats = nullptr;
};
std::string float_to_string(float v);     // std::to_string(float)
std::string float_ostream(float v);       // operator<<(ostream&, float) at default precision

}  // namespace orfx
