/* Stub (ours): see libheif/heif.h. */
#ifndef UHDR_B200_STUB_GMM_H
#define UHDR_B200_STUB_GMM_H
struct GainMapMetadata;
#endif
