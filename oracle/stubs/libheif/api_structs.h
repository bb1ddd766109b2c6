/* Stub (ours): intentionally empty, see libheif/heif.h. */
