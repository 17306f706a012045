// Genetic code with IUPAC ambiguity codes: the 180 (codon -> amino acid) facts of the reference's
// `codon_map` (src/common.hpp:279).  Stop = '.'.  Any other triplet, and any triplet shorter than 3
// bases, translates to '\0' there (std::map::operator[] on a missing key, common.hpp:431,492).
// Stored as 4-character strings "NNNa"; regenerate with tools/gen_codon_table.py.
#pragma once
static const char* const ORFX_CODON_TABLE[] = {
    "AAAK", "AACN", "AAGK", "AARK", "AATN", "AAYN", "ACAT", "ACBT", "ACCT", "ACDT", "ACGT", "ACHT",
    "ACKT", "ACMT", "ACNT", "ACRT", "ACST", "ACTT", "ACVT", "ACWT", "ACYT", "AGAR", "AGCS", "AGGR",
    "AGRR", "AGTS", "AGYS", "ATAI", "ATCI", "ATGM", "ATHI", "ATMI", "ATTI", "ATWI", "ATYI", "CAAQ",
    "CACH", "CAGQ", "CARQ", "CATH", "CAYH", "CCAP", "CCBP", "CCCP", "CCDP", "CCGP", "CCHP", "CCKP",
    "CCMP", "CCNP", "CCRP", "CCSP", "CCTP", "CCVP", "CCWP", "CCYP", "CGAR", "CGBR", "CGCR", "CGDR",
    "CGGR", "CGHR", "CGKR", "CGMR", "CGNR", "CGRR", "CGSR", "CGTR", "CGVR", "CGWR", "CGYR", "CTAL",
    "CTBL", "CTCL", "CTDL", "CTGL", "CTHL", "CTKL", "CTML", "CTNL", "CTRL", "CTSL", "CTTL", "CTVL",
    "CTWL", "CTYL", "GAAE", "GACD", "GAGE", "GARE", "GATD", "GAYD", "GCAA", "GCBA", "GCCA", "GCDA",
    "GCGA", "GCHA", "GCKA", "GCMA", "GCNA", "GCRA", "GCSA", "GCTA", "GCVA", "GCWA", "GCYA", "GGAG",
    "GGBG", "GGCG", "GGDG", "GGGG", "GGHG", "GGKG", "GGMG", "GGNG", "GGRG", "GGSG", "GGTG", "GGVG",
    "GGWG", "GGYG", "GTAV", "GTBV", "GTCV", "GTDV", "GTGV", "GTHV", "GTKV", "GTMV", "GTNV", "GTRV",
    "GTSV", "GTTV", "GTVV", "GTWV", "GTYV", "MGAR", "MGGR", "MGRR", "NNNX", "RAYB", "SARZ", "TAA.",
    "TACY", "TAG.", "TAR.", "TATY", "TAYY", "TCAS", "TCBS", "TCCS", "TCDS", "TCGS", "TCHS", "TCKS",
    "TCMS", "TCNS", "TCRS", "TCSS", "TCTS", "TCVS", "TCWS", "TCYS", "TGA.", "TGCC", "TGGW", "TGTC",
    "TGYC", "TRA.", "TTAL", "TTCF", "TTGL", "TTRL", "TTTF", "TTYF", "XXXX", "YTAL", "YTGL", "YTRL"
};
static const int ORFX_CODON_TABLE_N = 180;
