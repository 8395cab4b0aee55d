"""TEST INFRASTRUCTURE - the slice of Biopython that extract_barcodes.py touches (SeqIO fastq)."""
