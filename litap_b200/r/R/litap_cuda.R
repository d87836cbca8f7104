# litap_cuda.R -- drop-in replacements for the four hot internals of LITAP, same names, same
# arguments, same returned columns (SURVEY.md 8b):
#
#   calc_ddir2(db, verbose)        replaces R/flow_01_calc_ddir.R:2-118   (callers R/flow_mapper.R:157,354)
#   calc_shed4(db, verbose)        replaces R/flow_02_calc_shed.R:10-58   (callers R/flow_mapper.R:180,377)
#   calc_form(db, grid, verbose)   replaces R/form_01_calc_form.R:4-67    (caller  R/form_mapper.R:103)
#   calc_weti2(db, grid, verbose)  replaces R/form_02b_calc_weti.R:3-110  (caller  R/form_mapper.R:127)
#   slope_gc(db, grid)             replaces R/flow_05_slope_gc.R:73-164   (caller  R/flow_mapper.R:307)
#   calc_relz_length(db, idb, ...) replaces R/form_03_calc_relz.R:3-48 + R/form_04_calc_length.R:3-35 (callers R/form_mapper.R:154,177)
#
# Every function calls the CUDA library through .Call unconditionally: there is no CPU fallback
# (an error from the library, e.g. "no CUDA device", surfaces as an R error).  flow_mapper() and
# form_mapper() keep their signatures and their output column layout; pit removal (flow_03/04),
# the I/O stays the reference's R code.
#
# `db` is the reference's buffered tibble: one row per cell of the (nrow+2) x (ncol+2) grid,
# sorted by seqno (R/LITAP_load.R:273-332).

litap_dims <- function(db) {
  nr <- max(db$row); nc <- max(db$col)
  if (nrow(db) != nr * nc || is.unsorted(db$seqno))
    stop("litap: db must hold the full buffered grid sorted by seqno", call. = FALSE)
  c(nr, nc)
}

calc_ddir2 <- function(db, verbose = FALSE) {
  d <- litap_dims(db)
  if (verbose) message("  Calculating directions on the GPU")
  r <- .Call("litap_R_ddir", as.double(db$elev), d[1], d[2], FALSE)
  db$ddir <- r$ddir
  db$flatcell <- r$flatcell
  db$drec <- r$drec
  db
}

calc_shed4 <- function(db, verbose = FALSE) {
  d <- litap_dims(db)
  if (verbose) message("  Labelling watersheds and accumulating upslope cells on the GPU")
  r <- .Call("litap_R_shed", as.double(db$elev), db$ddir, d[1], d[2])
  db$initial_shed <- r$initial_shed
  db$upslope <- r$upslope
  db$ridge <- r$ridge
  # side columns the reference's calc_upslopes() leaves in its result
  # (R/flow_02_calc_shed.R:62-68); not saved (cols_order) but part of its own snapshot test
  # group_by(shedno) numbers the NA group as well (cells without a shed count 1, 2, ... among themselves)
  grp <- ifelse(is.na(db$initial_shed), 0, db$initial_shed)
  db$seqno_shed <- as.integer(stats::ave(db$seqno, grp, FUN = seq_along))
  db$drec_shed <- db$seqno_shed[db$drec]
  db
}

# calc_upslopes(db, type = "uced") (R/flow_02_calc_shed.R:60-79): what pit removal calls at
# R/flow_04_remove_pit_fast.R:73,345.  type = "upslope" is part of calc_shed4() above.
calc_upslopes <- function(db, type = c("upslope", "uced")) {
  type <- match.arg(type)
  if (type != "uced") stop("calc_upslopes(type = \"upslope\") is computed by calc_shed4()", call. = FALSE)
  d <- litap_dims(db)
  db$uced <- .Call("litap_R_uced", as.double(db$elev), db$ddir, db$shedno, d[1], d[2])
  db
}

# pit_stat1() (R/flow_03_pit_stat_fast.R:1-162): pour point and pit statistics of every shed
pit_stat1 <- function(db, w = NULL, shed = "shedno", method, verbose = FALSE) {
  d <- litap_dims(db)
  if (verbose) message("    Assessing pour points on the GPU")
  r <- .Call("litap_R_pit_stat", as.double(db$elev), db$ddir, db[[shed]], db$upslope, db$ridge, db$edge_map,
             d[1], d[2], as.integer(method == "landmapr"))
  stats <- tibble::as_tibble(r)
  stats$edge_pit <- stats$edge_pit == 1
  if (!is.null(w)) stats <- stats[stats$shedno %in% w, ]
  if ("pond_shed" %in% names(db))
    stats <- dplyr::left_join(stats, dplyr::distinct(dplyr::select(db, "shedno", "pond_shed")), by = "shedno")
  stats
}

calc_form <- function(db, grid, verbose = FALSE) {
  d <- litap_dims(db)
  if (verbose) message("  Calculating slope, aspect and curvature on the GPU")
  r <- .Call("litap_R_form", as.double(db$elev), d[1], d[2], as.double(grid))
  tibble::tibble(seqno = db$seqno, row = db$row, col = db$col,
                 slope_pct = r$slope_pct, slope_deg = r$slope_deg, aspect = r$aspect,
                 prof = r$prof, plan = r$plan, buffer = db$buffer)
}

calc_weti2 <- function(db, grid, verbose = FALSE) {
  d <- litap_dims(db)
  if (verbose) message("  Calculating wetness indices on the GPU")
  r <- .Call("litap_R_weti", as.double(db$elev), db$ddir, db$upslope, d[1], d[2], as.double(grid))
  db$qarea1 <- r$qarea1
  db$qarea2 <- r$qarea2
  db$qweti1 <- r$qweti1
  db$qweti2 <- r$qweti2
  db
}

slope_gc <- function(db, grid = 1) {
  if (!"buffer" %in% names(db)) db <- add_buffer(db)   # as the reference does (R/flow_05_slope_gc.R:75)
  d <- litap_dims(db)
  r <- .Call("litap_R_slope_gc", as.double(db$elev), d[1], d[2], as.double(grid))
  for (n in names(r)) db[[n]] <- r[[n]]                # same twelve columns, same order as the reference appends them
  db
}

# calc_relz() and calc_length() in one call: form_mapper() (R/form_mapper.R:154,177) may replace
#   db_relz <- calc_relz(db, idb, str_val, ridge_val, pond, verbose); db_length <- calc_length(db, db_relz, grid, verbose)
# by db_length <- calc_relz_length(db, idb, str_val, ridge_val, pond, grid); db_relz is db_length without the
# eight length columns (l2pit .. pstr2divl).  db / idb are the buffered tibbles form_mapper holds.
calc_relz_length <- function(db, idb, str_val, ridge_val, pond = NULL, grid = 1, verbose = FALSE) {
  d <- litap_dims(db)
  if (verbose) message("  Calculating relief and slope lengths on the GPU")
  if (is.null(pond) || nrow(pond) == 0)
    pond <- data.frame(shedno = integer(0), pit_row = integer(0), pit_col = integer(0), pit_elev = double(0))
  r <- .Call("litap_R_relz", d[1], d[2],
             list(as.double(db$elev), db$ddir, db$drec, db$upslope, db$fill_shed, db$local_shed),
             list(as.double(idb$elev), idb$ddir, idb$drec, idb$upslope, idb$inv_local_shed),
             list(pond$shedno, pond$pit_row, pond$pit_col, pond$pit_elev),
             as.double(str_val), as.double(ridge_val), as.double(grid))
  out <- db[, c("seqno", "elev", "row", "col", "buffer", "fill_shed")]
  for (n in names(r)) out[[n]] <- r[[n]]
  out
}

# The second traversal of flow_mapper() (R/flow_mapper.R:343-377) may skip the R-side invert()
# round trip for the direction step: calc_ddir2_inverted(db) == calc_ddir2(invert(db)) with the
# subtraction max(elev) - elev done on the device.  invert() itself (R/flow_06_invert.R) is still
# called by flow_mapper() because the inverted elevations are saved with dem_inverted.
calc_ddir2_inverted <- function(db, verbose = FALSE) {
  d <- litap_dims(db)
  r <- .Call("litap_R_ddir", as.double(db$elev), d[1], d[2], TRUE)
  db$elev <- max(db$elev, na.rm = TRUE) - db$elev
  db$ddir <- r$ddir
  db$flatcell <- r$flatcell
  db$drec <- r$drec
  db
}

.onLoad <- function(libname, pkgname) {
  dev <- as.integer(Sys.getenv("LITAP_CUDA_DEVICE", "-1"))
  .Call("litap_R_init", dev)
}
.onUnload <- function(libpath) .Call("litap_R_shutdown")
