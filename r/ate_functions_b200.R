# ate_functions_b200.R -- replacement bodies for the hot-path closures of ate_functions.R.
# Same names, same formals (new ones appended with defaults), same return values.
# NOT RUN IN THIS IMAGE (no R).  Source after ate_functions.R:  source("ate_functions_b200.R")
.ate_ctx <- NULL
ate_ctx <- function(device = 0L) {
  if (is.null(.ate_ctx)) .ate_ctx <<- .Call("C_ate_ctx_create", as.integer(device))
  .ate_ctx
}

.cv_glmnet_b200 <- function(x, y, family = 0L, penalty.factor = NULL, nfolds = 10L, foldid = NULL) {
  # cv.glmnet draws foldid = sample(rep(seq(nfolds), length = N)) from R's RNG; do it here so the
  # global stream advances exactly as with glmnet, then hand the ids down (validation mode)
  if (is.null(foldid)) foldid <- sample(rep(seq(nfolds), length = nrow(x)))
  storage.mode(x) <- "double"            # as.matrix() of all-integer covariates is an integer matrix
  if (!is.null(penalty.factor)) penalty.factor <- as.double(penalty.factor)
  m <- .Call("C_ate_cv_glmnet", ate_ctx(), as.integer(family), x, as.double(y), penalty.factor,
             as.integer(max(foldid)), as.integer(foldid), 0L)
  k <- seq_len(m$lmu)                    # the buffers hold nlambda = 100 slots; the path may stop earlier
  m$lambda <- m$lambda[k]; m$a0 <- m$a0[k]; m$beta <- m$beta[, k, drop = FALSE]
  m$cvm <- m$cvm[k]; m$cvsd <- m$cvsd[k]; m$dev.ratio <- m$dev.ratio[k]
  m
}

# ate_functions.R:89-108
ate_condmean_lasso <- function(dataset, treatment_var, outcome_var) {
  regs <- c(covariates, treatment_var)
  x <- as.matrix(dataset[regs]); y <- as.matrix(dataset[, outcome_var])
  pfac <- c(rep(1, length(covariates)), 0)
  model <- .cv_glmnet_b200(x, y, 0L, pfac)
  betaw <- model$beta[length(regs), model$idx.1se]          # coef(model)[treatment_var, ], s = "lambda.1se"
  data.frame(Method = "Single-equation LASSO", ATE = betaw, lower_ci = betaw, upper_ci = betaw)
}

# ate_functions.R:111-130
ate_lasso <- function(dataset, treatment_var, outcome_var) {
  regs <- c(covariates, treatment_var)
  x <- as.matrix(dataset[regs]); y <- as.matrix(dataset[, outcome_var])
  model <- .cv_glmnet_b200(x, y, 0L, rep(1, length(regs)))
  betaw <- model$beta[length(regs), model$idx.1se]
  data.frame(Method = "Usual LASSO", ATE = betaw, lower_ci = betaw, upper_ci = betaw)
}

# ate_functions.R:133-146
prop_score_lasso <- function(dataset, treatment_var) {
  x <- as.matrix(dataset[covariates]); w <- as.matrix(dataset[, treatment_var])
  model <- .cv_glmnet_b200(x, w, 1L)
  s <- model$idx.1se
  storage.mode(x) <- "double"
  .Call("C_ate_predict", ate_ctx(), x, model$a0[s], as.double(model$beta[, s]), TRUE)   # n x 1 matrix
}

# ate_functions.R:267-283
tau_hat_dr_est <- function(w, y, p, tauhat0x, tauhat1x) {
  bootstrap_sample_idx <- sample(length(w), length(w), replace = T)
  .Call("C_ate_tau_hat_dr_est", ate_ctx(), as.double(w), as.double(y), as.double(p),
        as.double(tauhat0x), as.double(tauhat1x), as.integer(bootstrap_sample_idx))
}

# ate_functions.R:241-263 (the tail of doubly_robust / doubly_robust_glm; nuisances as today)
# boot_idx (validation mode): integer matrix n x B, ONE REPLICATE PER COLUMN -- column b is the
# sample(n, n, replace = T) of replicate b, i.e. cbind() of the draws in loop order; values in 1..n
.dr_tail <- function(w, y, p, tauhat0x, tauhat1x, bootstrap_se, method, B = 1000, boot_idx = NULL) {
  if (!is.null(boot_idx)) { d <- dim(boot_idx); boot_idx <- as.integer(boot_idx); dim(boot_idx) <- d }
  ts <- .Call("C_ate_aipw_score", ate_ctx(), as.double(w), as.double(y), as.double(p),
              as.double(tauhat0x), as.double(tauhat1x))
  tau_hat <- ts[1]
  se_hat <- if (bootstrap_se == T) {
    attr(.Call("C_ate_bootstrap", ate_ctx(), as.double(w), as.double(y), as.double(p), as.double(tauhat0x),
               as.double(tauhat1x), as.integer(B), boot_idx, 1991), "sd")
  } else ts[2]
  data.frame(Method = method, ATE = tau_hat, lower_ci = tau_hat - se_hat * 1.96, upper_ci = tau_hat + se_hat * 1.96)
}

# coef(glmnet.fit, s = s): linear interpolation in lambda (glmnet's lambda.interp)
.coef_at <- function(model, s) {
  lam <- model$lambda; k <- length(lam)
  if (s >= lam[1]) return(model$beta[, 1]); if (s <= lam[k]) return(model$beta[, k])
  sfrac <- (lam[1] - s) / (lam[1] - lam[k]); lamf <- (lam[1] - lam) / (lam[1] - lam[k])
  coord <- approx(lamf, seq(lam), sfrac)$y; left <- floor(coord); right <- ceiling(coord)
  if (left == right) return(model$beta[, left])
  f <- (sfrac - lamf[right]) / (lamf[left] - lamf[right])
  model$beta[, left] * f + model$beta[, right] * (1 - f)
}

# lm() %>% summary() %>% coef() for the regression estimators: rows named like R names them
.lm_b200 <- function(x, y, weights = NULL, resid = FALSE) {
  storage.mode(x) <- "double"
  if (!is.null(weights)) weights <- as.double(weights)
  tab <- .Call("C_ate_lm", ate_ctx(), x, as.double(y), weights, resid)
  dimnames(tab) <- list(c("(Intercept)", colnames(x)), c("Estimate", "Std. Error"))
  tab
}
.ci_row <- function(method, tau_hat, se_hat)
  data.frame(Method = method, ATE = tau_hat, lower_ci = tau_hat - se_hat * 1.96, upper_ci = tau_hat + se_hat * 1.96)

# ate_functions.R:24-39
ate_condmean_ols <- function(dataset, treatment_var, outcome_var, method = "Direct Method") {
  x <- model.matrix(~ ., dataset[, names(dataset) != outcome_var])[, -1, drop = FALSE]
  reg_coef <- .lm_b200(x, dataset[, outcome_var])
  .ci_row(method, reg_coef[treatment_var, "Estimate"], reg_coef[treatment_var, "Std. Error"])
}

# ate_functions.R:44-64
prop_score_weight <- function(dataset, p, treatment_var, outcome_var, covariates, method = "Propensity_Weighting") {
  w <- dataset[, treatment_var]; y <- dataset[, outcome_var]
  tau_i <- ((w - p) * y) / (p * (1 - p)); ps_er <- w - p
  d <- as.matrix(dataset[, covariates] * ps_er)
  rss <- attr(.lm_b200(d, tau_i), "rss")                       # e <- residuals(summary(lm(tau_hat ~ .)))
  .ci_row(method, mean(tau_i), sqrt(rss / NROW(dataset)) / sqrt(NROW(dataset)))
}

# ate_functions.R:67-86
prop_score_ols <- function(dataset, p, treatment_var, outcome_var, method = "Propensity_Regression") {
  w <- dataset[, treatment_var]
  weights <- (w / p) + ((1 - w) / (1 - p))
  x <- matrix(w, ncol = 1, dimnames = list(NULL, treatment_var))
  reg_coef <- .lm_b200(x, dataset[, outcome_var], as.double(weights))
  .ci_row(method, reg_coef[treatment_var, "Estimate"], reg_coef[treatment_var, "Std. Error"])
}

# ate_functions.R:286-328 -- expansion, both cv.glmnet fits and the post-lasso OLS on the library
belloni <- function(dataset, treatment_var, outcome_var, method = "Belloni et.al") {
  x0 <- as.matrix(dataset[, covariates]); storage.mode(x0) <- "double"
  x <- .Call("C_ate_expand_deg2", ate_ctx(), x0)                                          # :288-296 (mains, then products)
  w <- as.matrix(dataset[treatment_var]); y <- as.matrix(dataset[outcome_var])
  model_xw <- .cv_glmnet_b200(x, w, 0L); model_xy <- .cv_glmnet_b200(x, y, 0L)            # :304-305
  s <- model_xw$lambda[model_xw$idx.min]
  c_xw <- model_xw$beta[, model_xw$idx.min]; c_xy <- .coef_at(model_xy, s)                # :308-309
  c_union <- unique(c(which(c_xw > 0), which(c_xy > 0))) - 1                              # :312-314
  x_restricted <- cbind(x[, c_union], w); colnames(x_restricted) <- NULL                 # :317
  reg_coef <- .lm_b200(x_restricted, y)                                                   # :320
  k <- nrow(reg_coef)                                                                     # the "x_restrictedW" row
  .ci_row(method, reg_coef[k, "Estimate"], reg_coef[k, "Std. Error"])
}

# ate_functions.R:211-264 -- same formals; only the two glm() calls and the tail change
doubly_robust_glm <- function(dataset, treatment_var, outcome_var, bootstrap_se = F) {
  mm <- function(d) model.matrix(~ ., d)[, -1, drop = FALSE]          # what glm builds for `~ .`
  plogis_lin <- function(x, b) .Call("C_ate_predict", ate_ctx(), x, b[1], as.double(b[-1]), TRUE)
  xo <- mm(dataset[, names(dataset) != outcome_var]); storage.mode(xo) <- "double"
  condmean <- .Call("C_ate_glm_binomial", ate_ctx(), xo, as.double(dataset[, outcome_var]))      # :218-220
  x1 <- xo; x1[, treatment_var] <- 1; x0 <- xo; x0[, treatment_var] <- 0                       # :221-226
  tauhat1x <- plogis_lin(x1, condmean); tauhat0x <- plogis_lin(x0, condmean)                   # :227-228
  xp <- xo[, colnames(xo) != treatment_var, drop = FALSE]
  pcoef <- .Call("C_ate_glm_binomial", ate_ctx(), xp, as.double(dataset[, treatment_var]))      # :231-233
  p <- plogis_lin(xp, pcoef)                                                                   # :234
  .dr_tail(dataset[, treatment_var], dataset[, outcome_var], p, tauhat0x, tauhat1x, bootstrap_se,
           "Doubly Robust with logistic regression PS")
}

# the composite of SURVEY.md section 8 (new; the reference's only sample-splitting precedent is :372-389)
crossfit_aipw <- function(dataset, treatment_var, outcome_var, bootstrap_se = T, B = 1000, K = 5,
                          seed = 1991, foldid = NULL, inner_ps = NULL, inner_out = NULL, boot_idx = NULL,
                          expand = 0L, n_continuous = 0L, shard = NULL) {
  # shard = c(rank, count): the rows of this ONE fit are shared by `count` R processes, one per GPU,
  # each of which called ate_join_ranks() first and passes the same dataset; all get the same row back
  x <- as.matrix(dataset[covariates]); storage.mode(x) <- "double"
  asint <- function(m) if (is.null(m)) NULL else { d <- dim(m); m <- as.integer(m); dim(m) <- d; m }
  foldid <- asint(foldid); inner_ps <- asint(inner_ps); inner_out <- asint(inner_out)
  boot_idx <- asint(boot_idx)            # n x B, one replicate per column (see .dr_tail)
  r <- .Call("C_ate_crossfit_aipw", ate_ctx(), x, as.double(dataset[, treatment_var]),
             as.double(dataset[, outcome_var]), as.integer(K), as.integer(if (bootstrap_se) B else 0L),
             as.integer(seed), foldid, inner_ps, inner_out, boot_idx, as.integer(expand), as.integer(n_continuous),
             if (is.null(shard)) NULL else as.integer(shard))
  se <- if (bootstrap_se) r[3] else r[2]
  data.frame(Method = "Cross-fitted AIPW with glmnet nuisances", ATE = r[1],
             lower_ci = r[1] - 1.96 * se, upper_ci = r[1] + 1.96 * se)
}

# Multi-GPU: one R process per GPU (mpirun -np G Rscript ...).  Rank 0 mints the NCCL id, Rmpi carries
# the 128 raw bytes, every process joins; the library then sums its per-round partials over the ranks
# on its own stream (include/ate_b200.h: ate_nccl_unique_id, ate_ctx_nccl_init).
ate_join_ranks <- function(rank = Rmpi::mpi.comm.rank(0), count = Rmpi::mpi.comm.size(0)) {
  id <- if (rank == 0) .Call("C_ate_nccl_unique_id") else raw(128)
  id <- Rmpi::mpi.bcast(id, type = 4, rank = 0, comm = 0)     # type 4: raw
  .Call("C_ate_ctx_nccl_init", ate_ctx(), id, as.integer(rank), as.integer(count))
  c(rank, count)
}
