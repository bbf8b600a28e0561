"""TEST INFRASTRUCTURE ONLY — CPU oracle for the PDF text framing row: the lines of
FileParser._parse_pdf_sync (app/file_parser.py:166-212) that come after the PDF engine, restated over
plain Python objects standing for what fitz returns (page.get_text(), table.extract(), page.get_images()).
Pinned by row 13 of the reference's .cache/42/cache.db (tests/golden/cache_rows.json): two pages of text
framed exactly this way."""
from __future__ import annotations


def frame_pdf(pages) -> str:
    """pages: sequence of (page_text, [table_data, ...], n_images)."""
    text_parts = []
    for page_num, (page_text, tables, n_images) in enumerate(pages):
        if page_text.strip():                                                     # :181-182
            text_parts.append(f"=== Page {page_num + 1} ===\n{page_text.strip()}")
        for table_num, table_data in enumerate(tables, 1):                        # :186-194
            if table_data:
                table_text = f"\n=== Table {table_num} on Page {page_num + 1} ===\n"
                for row in table_data:
                    if row and any(cell for cell in row if cell):
                        table_text += " | ".join(str(cell) if cell else "" for cell in row) + "\n"
                text_parts.append(table_text)
        if n_images:                                                              # :200-202
            text_parts.append(f"\n=== Images on Page {page_num + 1} ===\nFound {n_images} image(s)")
    if not text_parts:                                                            # :209-210
        return "PDF appears to be empty or contains no extractable text"
    return "\n\n".join(text_parts)                                                # :212
