"""Regenerates include/orfx_codon_table.h from the reference's codon_map literal
(src/common.hpp:279).  Only runs where /root/reference is mounted."""
import re, sys
ref = sys.argv[1] if len(sys.argv) > 1 else '/root/reference'
src = open(ref + '/src/common.hpp').read()
line = [l for l in src.splitlines() if 'codon_map' in l and '{{' in l][0]
pairs = sorted(re.findall(r'\{"([A-Z]{3})",\'(.)\'\}', line))
rows = ['    ' + ', '.join('"%s%s"' % (c, a) for c, a in pairs[i:i + 12]) for i in range(0, len(pairs), 12)]
print('\n'.join(r + ',' for r in rows))
print('// %d entries' % len(pairs))
