//! Native-endian u64 headers of the persistence format (same bytes as the reference's src/util.rs writes).
use std::fs::File;
use std::io::{Error, Read, Write};

pub fn read_u64(file: &mut File) -> Result<u64, Error> {
    let mut raw = [0u8; 8];
    file.read_exact(&mut raw)?;
    Ok(u64::from_ne_bytes(raw))
}

pub fn write_u64(file: &mut File, x: u64) -> Result<(), Error> {
    file.write_all(&x.to_ne_bytes())
}
