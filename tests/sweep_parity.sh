#!/bin/bash
# Byte-parity sweep: runs the reference binary (oracle/_ref/orfanage_ref) and a candidate CLI on the
# same seeded synthetic inputs over a grid of options and compares GTF, stats TSV and exit code.
# usage: sweep_parity.sh <candidate-cli> [workdir] [seeds...]
ROOT=$(cd "$(dirname "$0")/.." && pwd)
CLI=$(realpath ${1:-$ROOT/oracle/orfanage_oracle_cli})
WORK=${2:-/tmp/orfx_sweep}
shift 2 2>/dev/null
SEEDS=${@:-1 2 3}
REF=$ROOT/oracle/_ref/orfanage_ref
SYNTH=$ROOT/tools/orfx_synth
fail=0; n=0
run() { # dir, args...
  d=$1; shift
  (cd $d && $REF "$@" --output r.gtf --stats r.tsv $TPL >/dev/null 2>r.err; echo $? > r.rc)
  (cd $d && $CLI "$@" --output o.gtf --stats o.tsv $TPL >/dev/null 2>o.err; echo $? > o.rc)
  n=$((n+1))
  if ! cmp -s $d/r.gtf $d/o.gtf || ! cmp -s $d/r.tsv $d/o.tsv || ! cmp -s $d/r.rc $d/o.rc; then
    fail=$((fail+1)); echo "MISMATCH: $d $@ [tpl=$TPL] rc=$(cat $d/r.rc)/$(cat $d/o.rc)"; tail -2 $d/o.err
    diff $d/r.tsv $d/o.tsv | head -6
  fi
}
for seed in $SEEDS; do
  d=$WORK/s$seed; mkdir -p $d
  $SYNTH --out $d --seed $seed --loci 150 --chrom-len 3000000 --chroms 2 --queries 8 --templates 3 --template-files 3 --lower-frac 0.3 --iupac-rate 0.0005 --titin-frac 0.02 2>/dev/null
  TPL="template1.gtf template2.gtf template3.gtf"
  run $d --reference genome.fa --query query.gtf
  run $d --reference genome.fa --query query.gtf --mode BEST --cleanq
  run $d --reference genome.fa --query query.gtf --mode ALL
  run $d --reference genome.fa --query query.gtf --mode START_MATCH,FIRST
  run $d --reference genome.fa --query query.gtf --mode LONGEST --non_aug
  run $d --reference genome.fa --query query.gtf --mode LONGEST_MATCH,BEST --lpi 50 --ilpi 60 --mlpi 70 --minlen 300
  run $d --query query.gtf --mode LONGEST_MATCH,BEST
  run $d --query query.gtf --mode ALL --lpi 80
  run $d --reference genome.fa --query query.gtf --use_id
  run $d --reference genome.fa --query query.gtf --cleant
  run $d --reference genome.fa --query query.gtf --cleant --rescue --mode LONGEST_MATCH,BEST
  run $d --reference genome.fa --query query.gtf --cleant --rescue --non_aug --mode BEST
  run $d --reference genome.fa --query query.gtf --cleant --minlen 450 --use_id
  run $d --reference genome.fa --query query.gtf --overhang 30
  run $d --reference genome.fa --query query.gtf --overhang 300 --mode ALL
  run $d --reference genome.fa --query query.gtf --overhang 90 --spliced_overhang --mode BEST
  run $d --reference genome.fa --query query.gtf --overhang 2000 --spliced_overhang --non_aug --mode LONGEST
  run $d --reference genome.fa --query query.gtf --overhang 150 --cleant --rescue
  TPL="template1.gtf"
  run $d --reference genome.fa --query template2.gtf --keep_all_cds
  run $d --reference genome.fa --query template3.gtf --keep_cds_if_not_found --mode FIRST
  run $d --reference genome.fa --query template3.gtf --mode BEST
done
echo "runs=$n fail=$fail"
[ $fail -eq 0 ]
