# Step 2 of tools/r_pin/README.md.  Usage: Rscript tools/r_pin/make_golden.R <dir>
# Runs the reference's own calls on the exported inputs.  NOT run in the build image (no R there).
suppressMessages(library(glmnet))
dir <- commandArgs(trailingOnly = TRUE)[1]
wr <- function(x, name) write.table(format(as.matrix(x), digits = 17), file.path(dir, name), sep = ",",
                                    quote = FALSE, row.names = FALSE, col.names = FALSE)
for (tag in c("a", "b")) {
  x_raw <- as.matrix(read.csv(file.path(dir, paste0(tag, "_x_raw.csv")), header = FALSE))
  w <- scan(file.path(dir, paste0(tag, "_w.csv")), quiet = TRUE)
  y <- scan(file.path(dir, paste0(tag, "_y.csv")), quiet = TRUE)
  foldid <- scan(file.path(dir, paste0(tag, "_foldid.csv")), quiet = TRUE)
  ncont <- as.integer(strsplit(readLines(file.path(dir, paste0(tag, "_meta.txt")))[1], " ")[[1]][2])
  x <- x_raw
  x[, 1:ncont] <- scale(x_raw[, 1:ncont])                          # ate_replication.Rmd:72-74
  wr(x, paste0(tag, "_R_scaled.csv"))
  p0 <- ncol(x)
  xw <- cbind(x, w)
  pfac <- c(rep(1, p0), 0)                                         # ate_functions.R:98
  g <- glmnet(xw, y, alpha = 1, penalty.factor = pfac)             # :101-103 without the cv
  wr(rbind(g$lambda, g$a0, as.matrix(g$beta), g$dev.ratio), paste0(tag, "_R_gauss_path.csv"))
  cat(g$npasses, g$nulldev, "\n", file = file.path(dir, paste0(tag, "_R_gauss_scalars.txt")))
  b <- glmnet(x, w, alpha = 1, family = "binomial")                # :139-141 without the cv
  wr(rbind(b$lambda, b$a0, as.matrix(b$beta), b$dev.ratio), paste0(tag, "_R_binom_path.csv"))
  cat(b$npasses, b$nulldev, "\n", file = file.path(dir, paste0(tag, "_R_binom_scalars.txt")))
  cg <- cv.glmnet(xw, y, alpha = 1, penalty.factor = pfac, foldid = foldid)
  wr(rbind(cg$lambda, cg$cvm, cg$cvsd), paste0(tag, "_R_cv_gauss.csv"))
  cat(cg$lambda.min, cg$lambda.1se, "\n", file = file.path(dir, paste0(tag, "_R_cv_gauss_sel.txt")))
  cb <- cv.glmnet(x, w, alpha = 1, family = "binomial", foldid = foldid)
  wr(rbind(cb$lambda, cb$cvm, cb$cvsd), paste0(tag, "_R_cv_binom.csv"))
  cat(cb$lambda.min, cb$lambda.1se, "\n", file = file.path(dir, paste0(tag, "_R_cv_binom_sel.txt")))
  phat <- predict(cb, newx = x, type = "response")                 # :144 (s = "lambda.1se")
  wr(phat, paste0(tag, "_R_phat_1se.csv"))
  df <- data.frame(x, W = w, Y = y)
  gm <- glm(Y ~ ., family = binomial(link = "logit"), data = df)   # :218-220
  gp <- glm(W ~ ., family = binomial(link = "logit"), data = df[, names(df) != "Y"])  # :231-234
  wr(cbind(coef(gm)), paste0(tag, "_R_glm_outcome.csv")); wr(cbind(coef(gp)), paste0(tag, "_R_glm_ps.csv"))
  cat(gm$iter, gm$deviance, gp$iter, gp$deviance, "\n", file = file.path(dir, paste0(tag, "_R_glm_scalars.txt")))
  lmf <- summary(lm(Y ~ ., data = df))                             # :25-29
  wr(coef(lmf)[, 1:2], paste0(tag, "_R_lm.csv"))
}
set.seed(1991); s1 <- sample(50000, 20, replace = TRUE)             # :269 stream (current sample.kind)
suppressWarnings(RNGkind(sample.kind = "Rounding")); set.seed(1991); s0 <- sample(50000, 20, replace = TRUE)
wr(rbind(s1, s0), "R_sample_streams.csv")
cat(R.version.string, as.character(packageVersion("glmnet")), "\n", file = file.path(dir, "R_versions.txt"))
