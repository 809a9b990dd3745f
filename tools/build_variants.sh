#!/bin/bash
# Builds the staged kernel variants beside the measured default library (csrc/Makefile VARIANT / EXTRA); each is selected at run
# time with MRISR_B200_VARIANT=<name> and compared on a GPU with tools/ab_variant.sh <name>.
set -e
cd "$(dirname "$0")/../mri_superresolution_b200/csrc"
make -j"$(nproc)" VARIANT=incr EXTRA="-DMRISR_GNBWD_INCR"
make -j"$(nproc)" VARIANT=ups EXTRA="-DMRISR_UPSAMPLE_SLIDE -DMRISR_UPSAMPLE_BWD_SLIDE"
make -j"$(nproc)" VARIANT=dalpha EXTRA="-DMRISR_DALPHA_FUSED"
make -j"$(nproc)" VARIANT=all EXTRA="-DMRISR_GNBWD_INCR -DMRISR_UPSAMPLE_SLIDE -DMRISR_UPSAMPLE_BWD_SLIDE -DMRISR_DALPHA_FUSED"
ls -la ../lib
