/* Stub jerror.h (ours): intentionally empty, see stubs/jpeglib.h. */
