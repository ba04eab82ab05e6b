# Replacement for the body of Melissa:::melissa_inner (reference R/melissa_vb.R:232-529) that hands the
# whole VB run to libmelissa_b200.so.  Written without an R installation at hand (see INTEGRATION.md).
melissa_inner <- function(H, y, region_ind, cell_ind, K, basis, w, delta_0, alpha_0, beta_0,
                          vb_max_iter, epsilon_conv, is_parallel, no_cores, is_verbose, X,
                          n_gpus = getOption("melissa.n_gpus", 1L)) {
  # n_gpus: GPUs of this box to shard the cells over (the role no_cores had for mclapply, R/utils.R:30-43)
  res <- .Call(C_melissa_fit, X, as.integer(K), as.integer(basis$M), as.numeric(basis$mus), basis$gamma,
               w, delta_0, alpha_0, beta_0, as.integer(vb_max_iter), epsilon_conv, as.integer(n_gpus),
               isTRUE(is_verbose))
  M <- dim(w)[1]; D <- dim(w)[2]
  S <- array(res[[2]], dim = c(D, D, M, K))                         # block (k, m) at ((k-1)*M + m-1)*D*D
  W_Sigma <- lapply(seq_len(K), function(k) lapply(seq_len(M), function(m) S[, , m, k]))
  if (bitwAnd(res[[10]], 1L)) warning("Warning: Lower bound decreases!\n")   # :496
  if (bitwAnd(res[[10]], 2L)) warning("VB did not converge!\n")              # :512
  structure(list(W = res[[1]], W_Sigma = W_Sigma, r_nk = res[[3]], delta = res[[4]], alpha = res[[5]],
                 beta = res[[6]], basis = basis, pi_k = res[[7]], lb = res[[8]], delta_0 = delta_0,
                 alpha_0 = alpha_0, beta_0 = if (is.null(beta_0)) sqrt(res[[5]][1]) else beta_0),
            class = "melissa")
}
