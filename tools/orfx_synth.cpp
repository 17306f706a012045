// orfx_synth — seeded synthetic workload generator.
//
// Writes a random genome FASTA, one or more template GTFs (transcripts with CDS rows) and a
// query GTF (isoforms without CDS).  ORFs are planted into the genome (ATG + sense codons +
// stop, written along the spliced main isoform; reverse-complemented for '-' loci) so that
// the reference's rescue path (stop trim / extend, template-guided start choice) succeeds
// for most candidates, as it does on real annotation.  Everything is a pure function of the
// command line: the same flags give the same bytes on any machine (own PRNG, no libc rand).
//
// usage: orfx_synth --out DIR [--seed S] [--chroms N] [--chrom-len L] [--loci N]
//                   [--queries Q] [--templates T] [--template-files F] [--titin-frac X]
//                   [--lower-frac X] [--iupac-rate X] [--no-genome] [--line-width W]
//                   [--chrom-offset K] [--dot-frac X] [--edge-loci]
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>
#include <algorithm>

struct Rng {
    uint64_t s[4];
    static uint64_t splitmix(uint64_t& x){ uint64_t z=(x+=0x9E3779B97F4A7C15ULL); z=(z^(z>>30))*0xBF58476D1CE4E5B9ULL; z=(z^(z>>27))*0x94D049BB133111EBULL; return z^(z>>31); }
    explicit Rng(uint64_t seed){ for(int i=0;i<4;i++) s[i]=splitmix(seed); }
    static uint64_t rotl(uint64_t x,int k){ return (x<<k)|(x>>(64-k)); }
    uint64_t next(){ uint64_t r=rotl(s[1]*5,7)*9, t=s[1]<<17; s[2]^=s[0]; s[3]^=s[1]; s[1]^=s[2]; s[0]^=s[3]; s[2]^=t; s[3]=rotl(s[3],45); return r; }
    // uniform in [lo,hi]
    int64_t range(int64_t lo,int64_t hi){ if(hi<=lo) return lo; return lo+(int64_t)(next()%(uint64_t)(hi-lo+1)); }
    double unit(){ return (next()>>11)*(1.0/9007199254740992.0); }
    bool chance(double p){ return unit()<p; }
};

struct Seg { int64_t s,e; };            // 1-based inclusive genomic
typedef std::vector<Seg> Chain;

static int64_t chain_len(const Chain& c){ int64_t l=0; for(auto& x:c) l+=x.e-x.s+1; return l; }
static const char NT[4]={'A','C','G','T'};
static char comp(char c){ switch(c){case 'A':return 'T';case 'C':return 'G';case 'G':return 'C';case 'T':return 'A';} return c; }

// genomic coordinate of spliced offset `off` (5'->3') on chain c
static int64_t spliced2genome(const Chain& c,char strand,int64_t off){
    if(strand=='+'){ for(auto& x:c){ int64_t l=x.e-x.s+1; if(off<l) return x.s+off; off-=l; } return c.back().e; }
    for(int i=(int)c.size()-1;i>=0;i--){ int64_t l=c[i].e-c[i].s+1; if(off<l) return c[i].e-off; off-=l; }
    return c.front().s;
}
// spliced offset (5'->3') of genomic position g on chain c, or -1
static int64_t genome2spliced(const Chain& c,char strand,int64_t g){
    int64_t acc=0; bool found=false;
    for(auto& x:c){ if(g>=x.s&&g<=x.e){ acc+=g-x.s; found=true; break; } acc+=x.e-x.s+1; }
    if(!found) return -1;
    return strand=='+'?acc:chain_len(c)-acc-1;
}

struct Opts {
    std::string out=".";
    uint64_t seed=1;
    int chroms=1; int64_t chrom_len=2000000; int loci=20; int queries=10; int templates=2; int tfiles=1;
    double titin=0.01, lower=0.0, iupac=0.0, dot=0.0; bool genome=true, edge=false; int width=80; int chrom_offset=0;
    int64_t spacing=0;
};

static void write_fasta_seq(FILE* f,const std::string& name,const std::string& seq,int width){
    fprintf(f,">%s\n",name.c_str());
    size_t n=seq.size();
    std::string line; line.reserve(width+1);
    for(size_t i=0;i<n;i+=width){ size_t l=std::min((size_t)width,n-i); fwrite(seq.data()+i,1,l,f); fputc('\n',f); }
}

static void emit_tx(FILE* f,const std::string& chr,const char* src,const std::string& tid,const std::string& gid,char strand,
                    const Chain& ex,const Chain* cds,const std::vector<int>* phases,const std::string& extra){
    fprintf(f,"%s\t%s\ttranscript\t%lld\t%lld\t.\t%c\t.\ttranscript_id \"%s\"; gene_id \"%s\";%s\n",chr.c_str(),src,(long long)ex.front().s,(long long)ex.back().e,strand,tid.c_str(),gid.c_str(),extra.c_str());
    for(auto& x:ex) fprintf(f,"%s\t%s\texon\t%lld\t%lld\t.\t%c\t.\ttranscript_id \"%s\"; gene_id \"%s\";\n",chr.c_str(),src,(long long)x.s,(long long)x.e,strand,tid.c_str(),gid.c_str());
    if(cds) for(size_t i=0;i<cds->size();i++) fprintf(f,"%s\t%s\tCDS\t%lld\t%lld\t.\t%c\t%d\ttranscript_id \"%s\"; gene_id \"%s\";\n",chr.c_str(),src,(long long)(*cds)[i].s,(long long)(*cds)[i].e,strand,(*phases)[i],tid.c_str(),gid.c_str());
}

// clip exon chain to [lo,hi] genomic
static Chain clip(const Chain& ex,int64_t lo,int64_t hi){
    Chain r; for(auto& x:ex){ if(x.e<lo||x.s>hi) continue; r.push_back({std::max(x.s,lo),std::min(x.e,hi)}); } return r;
}
// GTF phase column for each CDS segment given the number of bases `lead` before the first full codon
static std::vector<int> gtf_phases(const Chain& cds,char strand,int lead){
    std::vector<int> ph(cds.size()); int64_t acc=0; // acc = coding bases consumed counted from the first codon boundary (negative lead)
    acc=-(int64_t)lead;
    if(strand=='+'){ for(size_t i=0;i<cds.size();i++){ ph[i]=(int)(((-acc)%3+3)%3); acc+=cds[i].e-cds[i].s+1; } }
    else { for(int i=(int)cds.size()-1;i>=0;i--){ ph[i]=(int)(((-acc)%3+3)%3); acc+=cds[i].e-cds[i].s+1; } }
    return ph;
}

static Chain mutate_isoform(Rng& r,const Chain& base,int64_t lo_lim,int64_t hi_lim){
    Chain c=base;
    int nops=(int)r.range(0,2);
    for(int op=0;op<nops;op++){
        int kind=(int)r.range(0,6);
        int n=(int)c.size();
        if(kind==0 && n>=3){ int i=(int)r.range(1,n-2); c.erase(c.begin()+i); }
        else if(kind==1 && n>=2){ // alt donor/acceptor
            int i=(int)r.range(0,n-1); int64_t d=r.range(1,30)*(r.chance(0.5)?1:-1);
            bool left=r.chance(0.5);
            if(left){ int64_t ns=c[i].s+d; int64_t lim=i>0?c[i-1].e+20:lo_lim; if(ns>=lim && ns<=c[i].e-9) c[i].s=ns; }
            else { int64_t ne=c[i].e+d; int64_t lim=i+1<n?c[i+1].s-20:hi_lim; if(ne<=lim && ne>=c[i].s+9) c[i].e=ne; }
        }
        else if(kind==2 && n>=2){ int i=(int)r.range(0,n-2); c[i].e=c[i+1].e; c.erase(c.begin()+i+1); } // intron retention
        else if(kind==3 && n>=2){ if(r.chance(0.5)) c.erase(c.begin()); else c.pop_back(); }
        else if(kind==4){ // shorten a terminal exon
            if(r.chance(0.5)){ int64_t l=c[0].e-c[0].s+1; if(l>20) c[0].s+=r.range(1,l-10); }
            else { int64_t l=c.back().e-c.back().s+1; if(l>20) c.back().e-=r.range(1,l-10); }
        }
        else if(kind==5 && n>=2){ // novel exon inside an intron
            int i=(int)r.range(0,n-2); int64_t gap=c[i+1].s-c[i].e-1;
            if(gap>160){ int64_t len=r.range(30,std::min<int64_t>(120,gap-60)); int64_t s=c[i].e+21+r.range(0,gap-len-41); c.insert(c.begin()+i+1,{s,s+len-1}); }
        }
        // kind 6 or inapplicable: unchanged
    }
    return c;
}

int main(int argc,char** argv){
    Opts o;
    for(int i=1;i<argc;i++){
        std::string a=argv[i];
        auto val=[&](){ if(i+1>=argc){ fprintf(stderr,"missing value for %s\n",a.c_str()); exit(2);} return std::string(argv[++i]); };
        if(a=="--out") o.out=val(); else if(a=="--seed") o.seed=strtoull(val().c_str(),0,10);
        else if(a=="--chroms") o.chroms=atoi(val().c_str()); else if(a=="--chrom-len") o.chrom_len=atoll(val().c_str());
        else if(a=="--loci") o.loci=atoi(val().c_str()); else if(a=="--queries") o.queries=atoi(val().c_str());
        else if(a=="--templates") o.templates=atoi(val().c_str()); else if(a=="--template-files") o.tfiles=atoi(val().c_str());
        else if(a=="--titin-frac") o.titin=atof(val().c_str()); else if(a=="--lower-frac") o.lower=atof(val().c_str());
        else if(a=="--iupac-rate") o.iupac=atof(val().c_str()); else if(a=="--no-genome") o.genome=false;
        else if(a=="--dot-frac") o.dot=atof(val().c_str()); else if(a=="--edge-loci") o.edge=true;
        else if(a=="--line-width") o.width=atoi(val().c_str()); else if(a=="--chrom-offset") o.chrom_offset=atoi(val().c_str());
        else { fprintf(stderr,"unknown option %s\n",a.c_str()); return 2; }
    }
    if(o.tfiles<1) o.tfiles=1;
    std::string fa_path=o.out+"/genome.fa", q_path=o.out+"/query.gtf";
    FILE* ffa=o.genome?fopen(fa_path.c_str(),"w"):nullptr;
    FILE* fq=fopen(q_path.c_str(),"w");
    std::vector<FILE*> ft;
    for(int k=0;k<o.tfiles;k++){ std::string p=o.out+"/template"+(o.tfiles>1?std::to_string(k+1):std::string(""))+".gtf"; ft.push_back(fopen(p.c_str(),"w")); if(!ft.back()){ perror(p.c_str()); return 1; } }
    if((o.genome&&!ffa)||!fq){ perror("open output"); return 1; }
    static const char* SRC[3]={"MANE","GENCODE","CHESS"};
    int per_chr=(o.loci+o.chroms-1)/o.chroms;
    int locus_id=0;
    long long n_q=0,n_t=0;
    for(int ci=0;ci<o.chroms;ci++){
        int cidx=ci+o.chrom_offset;
        std::string chr="chr"+std::to_string(cidx+1);
        Rng gr(o.seed*1000003ULL+ (uint64_t)cidx*7919ULL+17);   // genome stream
        Rng lr(o.seed*2000003ULL+ (uint64_t)cidx*104729ULL+29); // locus stream
        std::string seq;
        if(o.genome){
            seq.resize((size_t)o.chrom_len);
            for(int64_t i=0;i<o.chrom_len;i+=32){ uint64_t w=gr.next(); int64_t m=std::min<int64_t>(32,o.chrom_len-i); for(int64_t j=0;j<m;j++){ seq[(size_t)(i+j)]=NT[w&3]; w>>=2; } }
        }
        int nl=std::min(per_chr,o.loci-locus_id);
        if(nl<=0){ if(o.genome) write_fasta_seq(ffa,chr,seq,o.width); continue; }
        int64_t spacing=o.chrom_len/(nl+1);
        for(int li=0;li<nl;li++,locus_id++){
            bool titin=lr.chance(o.titin);
            char strand=lr.chance(0.5)?'+':'-';
            // --dot-frac: the locus is written with strand '.' (the reference reads such transcripts as reverse-strand
            // sequence); drawn only when asked for, so that existing seeds keep their streams
            const bool dot=o.dot>0 && lr.chance(o.dot);
            if(dot) strand='-';
            int nex=titin?(int)lr.range(100,360):(int)lr.range(4,14);
            int64_t budget=spacing-2000;      // locus must fit its slot
            // exon/intron lengths
            std::vector<int64_t> el(nex),il(nex-1);
            int64_t span=0;
            for(int i=0;i<nex;i++){ el[i]=lr.range(80,400); span+=el[i]; }
            int64_t imax=2500;
            for(int i=0;i<nex-1;i++){ il[i]=lr.range(100,imax); span+=il[i]; }
            if(span>budget){ // shrink introns proportionally (keep >=30)
                int64_t esum=0; for(auto x:el) esum+=x;
                double f=(double)std::max<int64_t>(budget-esum,30*(nex-1))/(double)(span-esum);
                span=esum; for(auto& x:il){ x=std::max<int64_t>(30,(int64_t)(x*f)); span+=x; }
            }
            int64_t slot_lo=(int64_t)li*spacing+1000, slot_hi=slot_lo+spacing-1000;
            if(span>slot_hi-slot_lo-10){ // does not fit: skip this locus deterministically
                continue;
            }
            int64_t start=slot_lo+lr.range(0,std::max<int64_t>(0,(slot_hi-slot_lo)-span-10));
            // --edge-loci: the first locus of a chromosome starts at its first base, the last one ends at its last base
            if(o.edge && li==0) start=1;
            else if(o.edge && li==nl-1) start=o.chrom_len-span+1;
            Chain ex(nex); int64_t p=start;
            for(int i=0;i<nex;i++){ ex[i].s=p; ex[i].e=p+el[i]-1; p+=el[i]; if(i<nex-1) p+=il[i]; }
            int64_t L=chain_len(ex);
            // plant ORF on the main isoform
            int64_t utr5=lr.range(20,std::min<int64_t>(400,L/4));
            int64_t max_codons=(L-utr5-30)/3; if(max_codons<20) max_codons=20;
            int64_t ncod=lr.range(std::max<int64_t>(20,max_codons/3),max_codons);   // incl. ATG and stop
            if(utr5+ncod*3>L){ ncod=(L-utr5)/3; }
            std::string orf; orf.reserve((size_t)ncod*3);
            orf+="ATG";
            static const char* STOPS[3]={"TAA","TAG","TGA"};
            for(int64_t c=1;c<ncod-1;c++){
                char cod[4]; for(;;){ uint64_t w=lr.next(); cod[0]=NT[w&3]; cod[1]=NT[(w>>2)&3]; cod[2]=NT[(w>>4)&3]; cod[3]=0;
                    if(!strcmp(cod,"TAA")||!strcmp(cod,"TAG")||!strcmp(cod,"TGA")) continue; break; }
                orf+=cod;
            }
            orf+=STOPS[lr.range(0,2)];
            // optional extra in-frame ATGs in the 5'UTR (alternative starts, no stop in between)
            int64_t up_clean=0;
            if(lr.chance(0.35)){ up_clean=std::min<int64_t>((utr5/3),lr.range(1,12)); }
            std::string up; // codons directly upstream of the ATG, in frame, stop-free
            for(int64_t c=0;c<up_clean;c++){
                if(lr.chance(0.3)) { up+="ATG"; continue; }
                char cod[4]; for(;;){ uint64_t w=lr.next(); cod[0]=NT[w&3]; cod[1]=NT[(w>>2)&3]; cod[2]=NT[(w>>4)&3]; cod[3]=0;
                    if(!strcmp(cod,"TAA")||!strcmp(cod,"TAG")||!strcmp(cod,"TGA")) continue; break; }
                up+=cod;
            }
            if(o.genome){
                std::string planted=up+orf; int64_t off0=utr5-(int64_t)up.size();
                for(size_t k=0;k<planted.size();k++){
                    int64_t g=spliced2genome(ex,strand,off0+(int64_t)k);
                    seq[(size_t)(g-1)]= strand=='+'?planted[k]:comp(planted[k]);
                }
            }
            int64_t cds5=spliced2genome(ex,strand,utr5), cds3=spliced2genome(ex,strand,utr5+ncod*3-1);
            int64_t clo=std::min(cds5,cds3), chi=std::max(cds5,cds3);
            std::string gid="G"+std::to_string(locus_id+1);
            bool have_t=!lr.chance(0.03), have_q=!lr.chance(0.03);
            // templates
            for(int fi=0;fi<o.tfiles;fi++){
                if(!have_t) break;
                int nt=(int)lr.range(1,std::max(1,o.templates));
                for(int k=0;k<nt;k++){
                    Chain tex=ex; int64_t tlo=clo,thi=chi; int lead=0;
                    int variant=k==0?0:(int)lr.range(0,5);
                    if(variant==1 && nex>=4){ // skip an internal exon (may break frame downstream: still a legal template)
                        int i=(int)lr.range(1,nex-2); tex.erase(tex.begin()+i);
                    } else if(variant==2){ // 5'-truncated CDS with a non-zero phase
                        int sh=(int)lr.range(1,40); int64_t np=utr5+sh; if(np<utr5+ncod*3-30){ int64_t g=spliced2genome(ex,strand,np); if(strand=='+') tlo=g; else thi=g; lead=(3-sh%3)%3; }
                    } else if(variant==3){ // CDS ends early (different stop position)
                        int64_t cut=lr.range(1,std::max<int64_t>(1,ncod/4))*3; int64_t np=utr5+ncod*3-1-cut; if(np>utr5+30){ int64_t g=spliced2genome(ex,strand,np); if(strand=='+') thi=g; else tlo=g; }
                    } else if(variant==4){ // alternative isoform of the template
                        tex=mutate_isoform(lr,ex,slot_lo,slot_hi);
                    }
                    Chain cds=clip(tex,tlo,thi);
                    if(cds.empty()) { tex=ex; cds=clip(ex,clo,chi); lead=0; }
                    std::vector<int> ph=gtf_phases(cds,strand,lead);
                    std::string tid=std::string(SRC[fi%3])+"."+gid+".T"+std::to_string(k+1);
                    std::string extra=" gene_name \"N"+std::to_string(locus_id+1)+"\"; Tag \"v"+std::to_string(variant)+"\";";
                    emit_tx(ft[fi],chr,SRC[fi%3],tid,gid,dot?'.':strand,tex,&cds,&ph,extra);
                    n_t++;
                }
            }
            // queries
            if(have_q){
                int nq=(int)lr.range(std::max(1,o.queries/2),std::max(1,o.queries+o.queries/2));
                for(int k=0;k<nq;k++){
                    Chain qex= k==0? ex : mutate_isoform(lr,ex,slot_lo,slot_hi);
                    std::string tid=gid+".Q"+std::to_string(k+1);
                    std::string extra=" gene_name \"N"+std::to_string(locus_id+1)+"\"; sample_count \""+std::to_string((int)lr.range(1,900))+"\";";
                    emit_tx(fq,chr,"StringTie",tid,gid,dot?'.':strand,qex,nullptr,nullptr,extra);
                    n_q++;
                }
            }
        }
        if(o.genome){
            if(o.lower>0||o.iupac>0){
                Rng mr(o.seed*3000017ULL+(uint64_t)cidx*31ULL+5);
                // soft-masked runs and sparse IUPAC codes
                if(o.lower>0){ size_t i=0; while(i<seq.size()){ size_t run=(size_t)mr.range(50,2000); if(mr.chance(o.lower)){ size_t e=std::min(seq.size(),i+run); for(size_t j=i;j<e;j++) seq[j]=(char)(seq[j]|0x20); } i+=run; } }
                if(o.iupac>0){ static const char IU[]="NRYKMSWBDHVnry"; size_t n=(size_t)(seq.size()*o.iupac); for(size_t k=0;k<n;k++){ size_t i=(size_t)mr.range(0,(int64_t)seq.size()-1); seq[i]=IU[mr.range(0,(int64_t)sizeof(IU)-2)]; } }
            }
            write_fasta_seq(ffa,chr,seq,o.width);
        }
    }
    if(ffa) fclose(ffa);
    fclose(fq); for(auto f:ft) fclose(f);
    fprintf(stderr,"orfx_synth: %d loci, %lld templates, %lld queries\n",locus_id,n_t,n_q);
    return 0;
}
