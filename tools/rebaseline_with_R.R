#!/usr/bin/env Rscript
# Pins the oracle to the reference itself, on the first machine that has R + BPRMeth + Melissa (SURVEY.md 8c).
#
# Runs the reference's un-exported Melissa:::melissa_inner (R/melissa_vb.R:232) on the two bundled data sets with exactly
# the initialisation the golden fixtures use (w0 / delta_0 written by tools/export_fixture_inputs.py as raw doubles), and
# dumps lb, r_nk and W as raw little-endian doubles next to the fixtures.  tests/test_r_dump.py picks the dumps up when
# they exist and compares the oracle (and, on a GPU, the CUDA path) with them; until then it reports "parity unpinned".
#
#   python tools/export_fixture_inputs.py            # writes tests/golden/r_inputs/*.bin  (needs no R)
#   Rscript tools/rebaseline_with_R.R                # writes tests/golden/r_dump/*.bin    (needs R, BPRMeth, Melissa)
suppressPackageStartupMessages({ library(BPRMeth); library(Melissa) })
root <- normalizePath(file.path(dirname(sub("--file=", "", grep("--file=", commandArgs(FALSE), value = TRUE))), ".."))
inp <- file.path(root, "tests", "golden", "r_inputs")
out <- file.path(root, "tests", "golden", "r_dump")
dir.create(out, showWarnings = FALSE, recursive = TRUE)
rd <- function(name, n) readBin(file.path(inp, name), "double", n = n, size = 8, endian = "little")
wr <- function(x, name) writeBin(as.double(x), file.path(out, name), size = 8, endian = "little")

run_one <- function(tag, X, basis, K, iters) {
  M <- length(X[[1]]); D <- basis$M + 1; N <- length(X)
  w <- array(rd(paste0(tag, "_w0.bin"), M * D * K), dim = c(M, D, K))        # column-major, as R stores it
  delta_0 <- rd(paste0(tag, "_delta_0.bin"), K)
  # the same preparation melissa() does before it calls melissa_inner (R/melissa_vb.R:101-131)
  X <- lapply(X, function(x) lapply(x, function(y) { if (is.matrix(y)) rownames(y) <- NULL; y }))
  region_ind <- lapply(X, function(x) which(!is.na(x)))
  cell_ind <- lapply(seq_len(M), function(m) which(!is.na(lapply(X, "[[", m))))
  H <- lapply(seq_len(N), function(n) Melissa:::init_design_matrix(basis = basis, X = X[[n]], coverage_ind = region_ind[[n]]))
  y <- lapply(seq_len(N), function(n) Melissa:::extract_y(X = X[[n]], coverage_ind = region_ind[[n]]))
  obj <- Melissa:::melissa_inner(H = H, y = y, region_ind = region_ind, cell_ind = cell_ind, K = K, basis = basis,
                                 w = w, delta_0 = delta_0, alpha_0 = .5, beta_0 = NULL, vb_max_iter = iters,
                                 epsilon_conv = 1e-5, is_parallel = FALSE, no_cores = 1, is_verbose = FALSE)
  wr(obj$lb, paste0(tag, "_lb.bin")); wr(obj$r_nk, paste0(tag, "_r_nk.bin")); wr(obj$W, paste0(tag, "_W.bin"))
  wr(c(obj$alpha, obj$beta, obj$delta, obj$pi_k), paste0(tag, "_scalars.bin"))
  cat(tag, ": lb =", tail(obj$lb, 1), " iterations with a bound:", length(obj$lb) - 1, "\n")
}

meta <- read.csv(file.path(inp, "runs.csv"), stringsAsFactors = FALSE)           # tag, dataset, K, iters, cpg mask file
for (i in seq_len(nrow(meta))) {
  dt <- if (meta$dataset[i] == "synth") melissa_synth_dt else melissa_encode_dt
  X <- dt$met
  if (nzchar(meta$mask[i])) {                                                    # the frozen 50 % CpG hold-out of config 2
    keep <- as.logical(readBin(file.path(inp, meta$mask[i]), "integer", n = meta$nnz[i], size = 1))
    pos <- 0
    X <- lapply(X, function(cell) lapply(cell, function(reg) {
      if (!is.matrix(reg)) return(reg)
      k <- keep[pos + seq_len(nrow(reg))]; pos <<- pos + nrow(reg)
      if (any(k)) reg[k, , drop = FALSE] else NA }))
  }
  run_one(meta$tag[i], X, dt$opts$basis_obj, meta$K[i], meta$iters[i])
}
