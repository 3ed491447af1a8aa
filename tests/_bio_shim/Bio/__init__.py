"""TEST-ONLY stand-in for the few Biopython (pinned ==1.83, Pipfile.lock:28-60) entry points
that /root/reference/kmerdb/{kmer,parse,graph,minimizer}.py import.  Biopython is absent from
this image; the k-mer arithmetic lives in the reference itself, Biopython only tokenises
FASTA/FASTQ and slices/reverse-complements strings.  Used ONLY by tests/golden/make_golden.py
to run the UNMODIFIED reference and freeze golden vectors.  Never imported by the product."""
__version__ = "1.83-shim"
