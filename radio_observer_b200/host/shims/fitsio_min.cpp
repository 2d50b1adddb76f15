/*
 * host/shims/fitsio_min.cpp -- see fitsio.h.  A streaming writer: header cards are collected until
 * the first pixel arrives, then the padded header goes to the file and every fits_write_pix call is
 * byte-swapped through a small buffer and written at its own file offset (without a seek when it
 * continues the previous one); closing pads the data unit to a whole 2880-byte block.  Nothing but the header is held in memory (the reference's
 * snapshots reach 100 MB).
 *
 * Cards follow the FITS fixed format the way CFITSIO fills them: keyword in columns 1-8, "= ",
 * numbers right-justified to column 30, strings quoted and padded to eight characters, real values
 * as %.7G (TFLOAT) / %.15G (TDOUBLE) with a decimal point forced in.
 */
#include "fitsio.h"

#include <cstdio>
#include <cstring>
#include <ctime>
#include <string>
#include <vector>

namespace {

const size_t BLOCK = 2880, CARD = 80;

std::string fixed(std::string s)
{
	s.resize(CARD, ' ');
	return s;
}

std::string with_comment(std::string body, const char *comment)
{
	if (comment && *comment) body.append(" / ").append(comment);
	return fixed(body);
}

std::string key_prefix(const char *key)
{
	char k[16];
	snprintf(k, sizeof k, "%-8.8s= ", key);
	return k;
}

std::string numeric_card(const char *key, const std::string &text, const char *comment)
{
	std::string v(text.size() < 20 ? 20 - text.size() : 0, ' ');
	return with_comment(key_prefix(key) + v + text, comment);
}

std::string real_text(double x, int digits)
{
	char b[64];
	snprintf(b, sizeof b, "%.*G", digits, x);
	const bool has_exp = strchr(b, 'E') != NULL, has_dot = strchr(b, '.') != NULL;
	if (has_exp && !has_dot) snprintf(b, sizeof b, "%.1E", x);
	else if (!has_exp && !has_dot) strcat(b, ".");
	return b;
}

}  // namespace

struct ro_min_fits {
	FILE *fp;
	std::string header;            // cards written so far
	bool header_on_disk;
	int bitpix;
	long width, height;
	size_t data_start;
	long   at;                     // file position after the last write, -1 = unknown
	std::vector<unsigned char> swap;

	bool flush_header()
	{
		if (header_on_disk) return true;
		std::string h = header + fixed("END");
		h.resize((h.size() + BLOCK - 1) / BLOCK * BLOCK, ' ');
		data_start = h.size();
		header_on_disk = true;
		at = (long)h.size();
		return fseek(fp, 0, SEEK_SET) == 0 && fwrite(h.data(), 1, h.size(), fp) == h.size();
	}
};

extern "C" int fits_create_file(fitsfile **fptr, const char *filename, int *status)
{
	if (*status > 0) return *status;
	if (!fptr || !filename) return *status = NULL_INPUT_PTR;
	std::string path(filename);
	const bool overwrite = !path.empty() && path[0] == '!';
	if (overwrite) path.erase(0, 1);
	const size_t bracket = path.find('[');          // "[compress]": see fitsio.h
	if (bracket != std::string::npos) path.erase(bracket);
	if (!overwrite) {
		if (FILE *probe = fopen(path.c_str(), "rb")) { fclose(probe); return *status = FILE_NOT_CREATED; }
	}
	FILE *fp = fopen(path.c_str(), "wb");
	if (!fp) return *status = FILE_NOT_CREATED;
	ro_min_fits *f = new ro_min_fits;
	f->fp = fp;
	f->header_on_disk = false;
	f->bitpix = 0;
	f->width = f->height = 0;
	f->data_start = 0;
	f->at = -1;
	*fptr = f;
	return *status;
}

extern "C" int fits_create_img(fitsfile *f, int bitpix, int naxis, long *naxes, int *status)
{
	if (*status > 0) return *status;
	if (!f || !naxes) return *status = NULL_INPUT_PTR;
	if (naxis != 2) return *status = BAD_DIMEN;
	f->bitpix = bitpix;
	f->width = naxes[0];
	f->height = naxes[1];
	f->header += numeric_card("SIMPLE", "T", "file does conform to FITS standard");
	f->header += numeric_card("BITPIX", std::to_string(bitpix), "number of bits per data pixel");
	f->header += numeric_card("NAXIS", "2", "number of data axes");
	f->header += numeric_card("NAXIS1", std::to_string(naxes[0]), "length of data axis 1");
	f->header += numeric_card("NAXIS2", std::to_string(naxes[1]), "length of data axis 2");
	f->header += numeric_card("EXTEND", "T", "FITS dataset may contain extensions");
	f->header += fixed("COMMENT   FITS (Flexible Image Transport System) format is defined in 'Astronomy");
	f->header += fixed("COMMENT   and Astrophysics', volume 376, page 359; bibcode: 2001A&A...376..359H");
	return *status;
}

extern "C" int fits_write_key(fitsfile *f, int datatype, const char *keyname, void *value,
                              const char *comment, int *status)
{
	if (*status > 0) return *status;
	if (!f || !keyname || !value) return *status = NULL_INPUT_PTR;
	if (f->header_on_disk) return *status = WRITE_ERROR;     // the header is closed once pixels follow
	switch (datatype) {
	case TSTRING: {
		std::string q("'");
		for (const char *c = (const char *)value; *c; c++) { q += *c; if (*c == '\'') q += '\''; }
		if (q.size() < 9) q.resize(9, ' ');
		q += '\'';
		if (q.size() < 20) q.resize(20, ' ');
		f->header += with_comment(key_prefix(keyname) + q, comment);
		break;
	}
	case TFLOAT:    f->header += numeric_card(keyname, real_text((double)*(const float *)value, 7), comment); break;
	case TDOUBLE:   f->header += numeric_card(keyname, real_text(*(const double *)value, 15), comment); break;
	case TSHORT:    f->header += numeric_card(keyname, std::to_string(*(const short *)value), comment); break;
	case TINT:      f->header += numeric_card(keyname, std::to_string(*(const int *)value), comment); break;
	case TLONG:     f->header += numeric_card(keyname, std::to_string(*(const long *)value), comment); break;
	case TLONGLONG: f->header += numeric_card(keyname, std::to_string(*(const long long *)value), comment); break;
	default: return *status = BAD_DATATYPE;
	}
	return *status;
}

extern "C" int fits_write_comment(fitsfile *f, const char *comment, int *status)
{
	if (*status > 0) return *status;
	if (!f) return *status = NULL_INPUT_PTR;
	if (f->header_on_disk) return *status = WRITE_ERROR;
	const std::string text(comment ? comment : "");
	size_t at = 0;
	do {
		f->header += fixed("COMMENT " + text.substr(at, 72));
		at += 72;
	} while (at < text.size());
	return *status;
}

extern "C" int fits_write_date(fitsfile *f, int *status)
{
	if (*status > 0) return *status;
	time_t now = time(NULL);
	struct tm utc;
	gmtime_r(&now, &utc);
	char stamp[32];
	strftime(stamp, sizeof stamp, "%Y-%m-%dT%H:%M:%S", &utc);
	return fits_write_key(f, TSTRING, "DATE", stamp, "file creation date (YYYY-MM-DDThh:mm:ss UT)", status);
}

extern "C" int fits_write_pix(fitsfile *f, int datatype, long *firstpix, long long nelem, void *array, int *status)
{
	if (*status > 0) return *status;
	if (!f || !firstpix || !array) return *status = NULL_INPUT_PTR;
	const int esz = (f->bitpix < 0 ? -f->bitpix : f->bitpix) / 8;
	if (!((datatype == TFLOAT && f->bitpix == FLOAT_IMG) || (datatype == TSHORT && f->bitpix == SHORT_IMG)))
		return *status = BAD_DATATYPE;
	const long long first = (firstpix[0] - 1) + (firstpix[1] - 1) * (long long)f->width;
	if (first < 0 || nelem < 0 || first + nelem > (long long)f->width * f->height) return *status = WRITE_ERROR;
	if (!f->flush_header()) return *status = WRITE_ERROR;
	const size_t bytes = (size_t)nelem * esz;
	if (f->swap.size() < bytes) f->swap.resize(bytes);
	const unsigned char *src = (const unsigned char *)array;
	unsigned char *dst = f->swap.data();
	if (esz == 4) {
		for (long long i = 0; i < nelem; i++) {
			dst[4 * i] = src[4 * i + 3]; dst[4 * i + 1] = src[4 * i + 2];
			dst[4 * i + 2] = src[4 * i + 1]; dst[4 * i + 3] = src[4 * i];
		}
	} else {
		for (long long i = 0; i < nelem; i++) { dst[2 * i] = src[2 * i + 1]; dst[2 * i + 1] = src[2 * i]; }
	}
	// the reference writes a raw-I/Q image one sample (8 bytes) per call, a million calls per file: a
	// write that continues where the last one ended goes to stdio's buffer without a seek (fseek flushes)
	const long where = (long)(f->data_start + (size_t)first * esz);
	if (where != f->at && fseek(f->fp, where, SEEK_SET) != 0) return *status = WRITE_ERROR;
	if (fwrite(f->swap.data(), 1, bytes, f->fp) != bytes) { f->at = -1; return *status = WRITE_ERROR; }
	f->at = where + (long)bytes;
	return *status;
}

extern "C" int fits_close_file(fitsfile *f, int *status)
{
	if (!f) return *status;
	bool ok = f->flush_header();
	const int esz = (f->bitpix < 0 ? -f->bitpix : f->bitpix) / 8;
	const size_t data = (size_t)esz * (size_t)f->width * (size_t)f->height;
	const size_t padded = (data + BLOCK - 1) / BLOCK * BLOCK;
	// rows never written stay zero; the data unit ends on a block boundary
	if (ok && fseek(f->fp, 0, SEEK_END) == 0) {
		const long end = ftell(f->fp);
		if (end >= 0 && (size_t)end < f->data_start + padded) {
			const std::vector<unsigned char> zeros(f->data_start + padded - (size_t)end, 0);
			ok = fwrite(zeros.data(), 1, zeros.size(), f->fp) == zeros.size();
		}
	} else {
		ok = false;
	}
	if (fclose(f->fp) != 0) ok = false;
	delete f;
	if (!ok && *status <= 0) *status = WRITE_ERROR;
	return *status;
}

extern "C" void fits_get_errstatus(int status, char *errtext)
{
	const char *what = status == FILE_NOT_CREATED ? "could not create file" :
	                   status == WRITE_ERROR ? "error writing to file" :
	                   status == BAD_DATATYPE ? "unsupported data type" :
	                   status == BAD_DIMEN ? "unsupported number of axes" :
	                   status == NULL_INPUT_PTR ? "null pointer argument" : "error";
	snprintf(errtext, FLEN_STATUS, "%.24s %d", what, status);
}

extern "C" int fits_read_errmsg(char *errmsg)
{
	if (errmsg) errmsg[0] = 0;
	return 0;                                    // no message stack
}
