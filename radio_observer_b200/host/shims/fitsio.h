/*
 * host/shims/fitsio.h -- the nine CFITSIO entry points the reference's FITSWriter uses
 * (src/FITSWriter.cpp:44-148: fits_create_file, fits_create_img, fits_write_key,
 * fits_write_comment, fits_write_date, fits_write_pix, fits_close_file; src/FITSWriter.h:31,39:
 * fits_get_errstatus, fits_read_errmsg), for building radio-observer-cuda on a machine without
 * CFITSIO.  fitsio_min.cpp behind it streams an uncompressed primary HDU straight to the file.
 * "[compress]" (tile compression, the reference's default) needs the real library: the suffix is
 * accepted and the file is written uncompressed.  host/Makefile uses the system's <fitsio.h> and
 * -lcfitsio instead whenever they are installed.
 */
#ifndef RO_HOST_FITSIO_MIN_H
#define RO_HOST_FITSIO_MIN_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ro_min_fits fitsfile;

enum { FLEN_FILENAME = 1025, FLEN_KEYWORD = 75, FLEN_CARD = 81, FLEN_VALUE = 71, FLEN_COMMENT = 73,
       FLEN_ERRMSG = 81, FLEN_STATUS = 31 };

/* image types (BITPIX) */
enum { BYTE_IMG = 8, SHORT_IMG = 16, LONG_IMG = 32, LONGLONG_IMG = 64, FLOAT_IMG = -32, DOUBLE_IMG = -64 };
/* value types of fits_write_key / fits_write_pix */
enum { TSTRING = 16, TSHORT = 21, TINT = 31, TLONG = 41, TFLOAT = 42, TLONGLONG = 81, TDOUBLE = 82 };
/* status codes */
enum { FILE_NOT_CREATED = 105, WRITE_ERROR = 106, NULL_INPUT_PTR = 115, BAD_DIMEN = 320, BAD_DATATYPE = 410 };

int  fits_create_file(fitsfile **fptr, const char *filename, int *status);
int  fits_create_img(fitsfile *fptr, int bitpix, int naxis, long *naxes, int *status);
int  fits_write_key(fitsfile *fptr, int datatype, const char *keyname, void *value, const char *comment, int *status);
int  fits_write_comment(fitsfile *fptr, const char *comment, int *status);
int  fits_write_date(fitsfile *fptr, int *status);
int  fits_write_pix(fitsfile *fptr, int datatype, long *firstpix, long long nelem, void *array, int *status);
int  fits_close_file(fitsfile *fptr, int *status);
void fits_get_errstatus(int status, char *errtext);
int  fits_read_errmsg(char *errmsg);

#ifdef __cplusplus
}
#endif

#endif
