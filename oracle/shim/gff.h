// TEST INFRASTRUCTURE ONLY (oracle). Minimal stand-in for the gclib surface that the
// reference's own sources (orfanage.cpp, orfcompare.cpp, src/Transcriptome.{h,cpp})
// link against.  gclib itself (gpertea/gclib @dcb0141c, CMakeLists.txt:58-66 of the
// reference) is fetched from the network by the reference's build and is not on this
// box, so only the symbols the reference calls are provided here (list: SURVEY.md
// §8c / Appendix B).  This file contains no ORFanage logic: it parses GTF rows into
// GffObj records and serves FASTA substrings.  It is never linked into the product.
#ifndef ORFX_ORACLE_GCLIB_SHIM_GFF_H
#define ORFX_ORACLE_GCLIB_SHIM_GFF_H

#include <cstdio>
#include <cstdint>
#include <cstring>
#include <climits>
#include <string>
#include <vector>
#include <map>
#include <sys/types.h>

#ifndef MAX_INT
#define MAX_INT INT_MAX
#endif

int Gstrcmp(const char* a, const char* b);
char ntComplement(char c);

template <class T> class ShimPtrList {
public:
    std::vector<T*> v;
    int Count() const { return (int)v.size(); }
    T* Get(int i) const { return v[i]; }
    T* operator[](int i) const { return v[i]; }
};

struct GffExon { uint start; uint end; };
struct GffAttr { int attr_id; char* attr_val; };

struct GffNameList {
    std::vector<std::string> names;
    std::map<std::string,int> ids;
    int addName(const std::string& n){
        auto it = ids.find(n);
        if(it!=ids.end()) return it->second;
        int id=(int)names.size(); names.push_back(n); ids[n]=id; return id;
    }
    const char* getName(int id){ return names[id].c_str(); }
};
struct GffNames { GffNameList attrs; };

class GffObj {
public:
    uint start=0, end=0;
    char strand='.';
    uint CDstart=0, CDend=0;
    char CDphase='.';
    ShimPtrList<GffExon> exons;
    ShimPtrList<GffAttr>* attrs=nullptr;
    GffNames* names=nullptr;

    bool hasCDS() const { return has_cds_; }
    bool isTranscript() const { return true; }
    const char* getID() const { return id_.c_str(); }
    char* getGeneID() { return has_gid_ ? const_cast<char*>(gid_.c_str()) : nullptr; }
    const char* getTrackName() const { return track_.c_str(); }
    const char* getGSeqName() const { return seq_.c_str(); }

    // shim-private
    std::string id_, gid_, track_, seq_;
    bool has_gid_=false, has_cds_=false, has_tx_line_=false;
    std::vector<std::pair<uint,uint> > cds_segs_;
    std::vector<char> cds_frames_;
    ~GffObj();
};

class GffReader {
public:
    ShimPtrList<GffObj> gflst;
    GffReader(FILE* f, bool t_only=false, bool sortbyloc=false);
    ~GffReader();
    void keepAttrs(bool keep=true){ (void)keep; }
    void readAll();
private:
    FILE* fh;
    GffNames* names;
};

class GFaSeqGet {
public:
    explicit GFaSeqGet(const std::string* s): seq(s) {}
    int getseqlen() const { return (int)seq->size(); }
    // 1-based start; returns a pointer to len raw bytes (case preserved)
    const char* subseq(uint cstart, int& clen);
private:
    const std::string* seq;
};

class GFastaDb {
public:
    GFastaDb() {}
    ~GFastaDb();
    void init(const char* fpath);
    std::map<std::string,std::string> seqs;
    std::map<std::string,GFaSeqGet*> getters;
};

GFaSeqGet* fastaSeqGet(GFastaDb& db, const char* name);

class GFastaIndex {
public:
    GFastaIndex(const char* fa, const char* fai){ (void)fa; (void)fai; }
    bool hasIndex(){ return true; }   // the stand-in serves sequences from memory; no .fai needed
    void buildIndex(){}
    int getCount(){ return 1; }
    int storeIndex(FILE* f){ (void)f; return 1; }
};

extern GFastaDb gfasta;   // defined in shim.cpp (the reference defines it in gff_utils.cpp:72, not compiled here)

#endif
