// TEST INFRASTRUCTURE ONLY (oracle). See gff.h in this directory.
#include "gff.h"
#include <stdexcept>
#include <algorithm>
#include <fstream>

GFastaDb gfasta;

int Gstrcmp(const char* a, const char* b){
    if(a==nullptr || b==nullptr) return a==b ? 0 : (a==nullptr ? -1 : 1);
    return strcmp(a,b);
}

// IUPAC complement, case preserving; bytes outside the IUPAC set are returned unchanged.
char ntComplement(char c){
    static const char* defs = "AaCcTtGgUuMmRrWwSsYyKkVvHhDdBbNnXx-*";
    static const char* comp = "TtGgAaCcAaKkYyWwSsRrMmBbDdHhVvNnXx-*";
    static char table[256];
    static bool init=false;
    if(!init){
        for(int i=0;i<256;i++) table[i]=(char)i;
        for(int i=0;defs[i];i++) table[(unsigned char)defs[i]]=comp[i];
        init=true;
    }
    return table[(unsigned char)c];
}

GffObj::~GffObj(){
    for(auto e: exons.v) delete e;
    if(attrs){
        for(auto a: attrs->v){ free(a->attr_val); delete a; }
        delete attrs;
    }
}

GffReader::GffReader(FILE* f,bool,bool): fh(f), names(new GffNames()) {}
GffReader::~GffReader(){
    for(auto o: gflst.v) delete o;
    delete names;
    if(fh) fclose(fh);
}

static void split_tab(const std::string& line, std::vector<std::string>& out){
    out.clear();
    size_t p=0;
    while(true){
        size_t q=line.find('\t',p);
        if(q==std::string::npos){ out.push_back(line.substr(p)); break; }
        out.push_back(line.substr(p,q-p));
        p=q+1;
    }
}

// parses `key "value"; key "value"; key value;` into ordered pairs
static void parse_attrs(const std::string& s, std::vector<std::pair<std::string,std::string> >& out){
    out.clear();
    size_t i=0, n=s.size();
    while(i<n){
        while(i<n && (s[i]==' '||s[i]==';'||s[i]=='\t')) i++;
        if(i>=n) break;
        size_t k0=i;
        while(i<n && s[i]!=' ' && s[i]!='\t' && s[i]!=';') i++;
        std::string key=s.substr(k0,i-k0);
        while(i<n && (s[i]==' '||s[i]=='\t')) i++;
        std::string val;
        if(i<n && s[i]=='"'){
            i++;
            size_t v0=i;
            while(i<n && s[i]!='"') i++;
            val=s.substr(v0,i-v0);
            if(i<n) i++;
        } else {
            size_t v0=i;
            while(i<n && s[i]!=';') i++;
            val=s.substr(v0,i-v0);
            while(!val.empty() && val.back()==' ') val.pop_back();
        }
        out.push_back(std::make_pair(key,val));
    }
}

void GffReader::readAll(){
    std::map<std::string,GffObj*> by_id;
    std::string line;
    std::vector<std::string> col;
    std::vector<std::pair<std::string,std::string> > av;
    char* buf=nullptr; size_t cap=0; ssize_t len;
    while((len=getline(&buf,&cap,fh))>=0){
        while(len>0 && (buf[len-1]=='\n'||buf[len-1]=='\r')) len--;
        if(len==0 || buf[0]=='#') continue;
        line.assign(buf,len);
        split_tab(line,col);
        if(col.size()<9) continue;
        const std::string& feat=col[2];
        bool is_tx = feat=="transcript" || feat=="mRNA";
        bool is_exon = feat=="exon";
        bool is_cds = feat=="CDS";
        if(!is_tx && !is_exon && !is_cds) continue;
        parse_attrs(col[8],av);
        std::string tid; bool has_gid=false; std::string gid;
        for(auto& kv: av){
            if(kv.first=="transcript_id" && tid.empty()) tid=kv.second;
            else if(kv.first=="gene_id" && !has_gid){ has_gid=true; gid=kv.second; }
        }
        if(tid.empty()) continue;
        GffObj* o;
        auto it=by_id.find(tid);
        if(it==by_id.end()){
            o=new GffObj();
            o->id_=tid; o->seq_=col[0]; o->track_=col[1];
            o->strand=col[6].empty()?'.':col[6][0];
            o->names=names;
            by_id[tid]=o;
            gflst.v.push_back(o);
        } else o=it->second;
        if(has_gid && !o->has_gid_){ o->has_gid_=true; o->gid_=gid; }
        uint s=(uint)strtoul(col[3].c_str(),nullptr,10);
        uint e=(uint)strtoul(col[4].c_str(),nullptr,10);
        if(s>e) std::swap(s,e);
        if(is_tx){
            o->has_tx_line_=true;
            o->track_=col[1];
            if(!o->attrs) o->attrs=new ShimPtrList<GffAttr>();
            for(auto& kv: av){
                if(kv.first=="transcript_id") continue;
                GffAttr* a=new GffAttr();
                a->attr_id=names->attrs.addName(kv.first);
                a->attr_val=strdup(kv.second.c_str());
                o->attrs->v.push_back(a);
            }
        } else if(is_exon){
            GffExon* ex=new GffExon(); ex->start=s; ex->end=e;
            o->exons.v.push_back(ex);
        } else {
            o->has_cds_=true;
            o->cds_segs_.push_back(std::make_pair(s,e));
            o->cds_frames_.push_back(col[7].empty()?'.':col[7][0]);
        }
    }
    free(buf);
    for(auto o: gflst.v){
        if(o->exons.v.empty()){ // CDS-only record: the coding segments act as exons
            for(auto& c: o->cds_segs_){ GffExon* ex=new GffExon(); ex->start=c.first; ex->end=c.second; o->exons.v.push_back(ex); }
        }
        std::sort(o->exons.v.begin(),o->exons.v.end(),[](GffExon* a,GffExon* b){ return a->start!=b->start ? a->start<b->start : a->end<b->end; });
        if(!o->exons.v.empty()){
            o->start=o->exons.v.front()->start;
            o->end=o->exons.v.back()->end;
        }
        if(o->has_cds_){
            uint mn=UINT_MAX,mx=0; size_t imn=0,imx=0;
            for(size_t i=0;i<o->cds_segs_.size();i++){
                if(o->cds_segs_[i].first<mn){ mn=o->cds_segs_[i].first; imn=i; }
                if(o->cds_segs_[i].second>mx){ mx=o->cds_segs_[i].second; imx=i; }
            }
            o->CDstart=mn; o->CDend=mx;
            o->CDphase = o->strand=='-' ? o->cds_frames_[imx] : o->cds_frames_[imn];
        }
    }
}

const char* GFaSeqGet::subseq(uint cstart,int& clen){
    if(cstart<1) cstart=1;
    if((size_t)(cstart-1)+(size_t)clen>seq->size()) clen=(int)(seq->size()-(cstart-1));
    return seq->data()+(cstart-1);
}

GFastaDb::~GFastaDb(){ for(auto& g: getters) delete g.second; }

void GFastaDb::init(const char* fpath){
    FILE* f=fopen(fpath,"r");
    if(!f) throw std::runtime_error(std::string("Cannot open FASTA: ")+fpath);
    char* buf=nullptr; size_t cap=0; ssize_t len;
    std::string* cur=nullptr;
    while((len=getline(&buf,&cap,f))>=0){
        while(len>0 && (buf[len-1]=='\n'||buf[len-1]=='\r')) len--;
        if(len==0) continue;
        if(buf[0]=='>'){
            ssize_t e=1;
            while(e<len && buf[e]!=' ' && buf[e]!='\t') e++;
            std::string name(buf+1,e-1);
            cur=&seqs[name];
            cur->clear();
        } else if(cur){
            cur->append(buf,len);
        }
    }
    free(buf);
    fclose(f);
    for(auto& kv: seqs) getters[kv.first]=new GFaSeqGet(&kv.second);
}

GFaSeqGet* fastaSeqGet(GFastaDb& db,const char* name){
    auto it=db.getters.find(name);
    return it==db.getters.end()?nullptr:it->second;
}
