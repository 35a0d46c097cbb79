"""Command-line shims with the reference's argparse surface (drop-in for
recipes/timit/local_pyspeech/make_{FDLPspectrum,melspectrum}_feats.sh via --src_dir)."""
