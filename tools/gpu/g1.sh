set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/g1_smi.txt
nproc >> gpurun_out/g1_smi.txt; lscpu | grep "Model name" >> gpurun_out/g1_smi.txt
bash tests/sweep_parity.sh orfanage_b200/bin/orfanage /tmp/sw 1 2 3 > gpurun_out/g1_sweep.log 2>&1
mkdir -p /tmp/w1 && tools/orfx_synth --out /tmp/w1 --seed 7 --loci 2000 --chroms 4 --chrom-len 50000000 --queries 12 --templates 4 --template-files 1 > gpurun_out/g1_synth.log 2>&1
cd /tmp/w1
( time /root/repo/orfanage_b200/bin/orfanage --reference genome.fa --query query.gtf --output o.gtf --stats o.tsv --mode LONGEST_MATCH,BEST --timing template.gtf ) > /root/repo/gpurun_out/g1_gpu.log 2>&1
( time /root/repo/oracle/_ref/orfanage_ref --reference genome.fa --query query.gtf --output r.gtf --stats r.tsv --mode LONGEST_MATCH,BEST --threads 1 template.gtf ) > /root/repo/gpurun_out/g1_ref1.log 2>&1
cmp o.gtf r.gtf >> /root/repo/gpurun_out/g1_gpu.log 2>&1; cmp o.tsv r.tsv >> /root/repo/gpurun_out/g1_gpu.log 2>&1
( time /root/repo/oracle/_ref/orfanage_ref --reference genome.fa --query query.gtf --output r8.gtf --mode LONGEST_MATCH,BEST --threads $(nproc) template.gtf ) > /root/repo/gpurun_out/g1_refN.log 2>&1
grep -c transcript o.gtf >> /root/repo/gpurun_out/g1_gpu.log
