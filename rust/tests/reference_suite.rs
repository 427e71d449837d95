//! The golden vectors of the reference's own tests (inputs and expected outputs as listed in
//! tests/golden/reference_tests.json, each entry citing the tests.rs lines it was transcribed from), run against this
//! facade.  Needs a B200 and liblinalg_b200.so:  LINALG_B200_DIR=../linear_algebra_b200 cargo test
//! NOT COMPILED IN THE BUILD IMAGE (no rustc there); tests/reference_suite.cpp runs the same cases through the C++ mirror.
use linear_algebra::matrix::{mul_column_row, mul_host, Matrix};
use linear_algebra::vector::{dot, mul_transpose_mat, Vector};
use linear_algebra::{create_kernel_from_source, traits::Parameter};

type TestType = u32;

#[test]
fn f32_vec_mat_mul() {
    let a = Vector::from_vec(vec![1.0f32, 2.0, 3.0, 4.0]);
    let m = Matrix::from_vec(vec![1.0f32, 0.0, 0.0, 0.0, 2.0, 0.0, 0.0, 0.0, 3.0, 1.0, 0.0, 0.0], 4, 3);
    assert_eq!((&a * &m).to_vec(), vec![5.0f32, 4.0, 9.0]);
}

#[test]
fn vec_mat_mul() {
    let a = Vector::<TestType>::from_vec(vec![1, 2, 3, 4]);
    let m = Matrix::<TestType>::from_vec(vec![1, 0, 0, 0, 2, 0, 0, 0, 3, 1, 0, 0], 4, 3);
    assert_eq!((&a * &m).to_vec(), vec![5, 4, 9]);
}

#[test]
fn vec_transpose_mat_mul() {
    let a = Vector::<TestType>::from_vec(vec![1, 2, 3, 4]);
    let m = Matrix::<TestType>::from_vec(vec![1, 0, 0, 1, 0, 2, 0, 0, 0, 0, 3, 0], 3, 4);
    assert_eq!(mul_transpose_mat(&a, &m).to_vec(), vec![5, 4, 9]);
}

#[test]
fn mat_mat_mul() {
    let a = Matrix::<TestType>::from_vec(vec![1, 2, 3, 4, 5, 6], 2, 3);
    let b = Matrix::<TestType>::from_vec(vec![7, 8, 9, 10, 11, 12], 3, 2);
    assert_eq!((&a * &b).to_vec(), vec![58, 64, 139, 154]);
    assert_eq!(mul_host::<TestType>(&[1, 2, 3, 4, 5, 6], &[7, 8, 9, 10, 11, 12], 2, 3, 2), vec![58, 64, 139, 154]);
}

#[test]
fn vec_sum() {
    for &(size, expect) in &[(1u32, 0u32), (2, 1), (8, 28), (16, 120), (64, 2016), (128, 8128), (971, 470935)] {
        let v: Vec<TestType> = (0..size).collect();
        assert_eq!(Vector::from_vec(v).sum(), expect);
    }
}

#[test]
fn vec_vec_dot() {
    let a: Vec<TestType> = (1..971).collect();
    let b: Vec<TestType> = (3..973).rev().collect();
    assert_eq!(dot(&Vector::from_vec(a), &Vector::from_vec(b)), 153524810);
}

#[test]
fn vec_add_and_scale() {
    let a = Vector::<TestType>::from_vec(vec![1, 2, 3, 4]);
    assert_eq!((&a + &a).to_vec(), vec![2, 4, 6, 8]);
    assert_eq!((&a * 2).to_vec(), vec![2, 4, 6, 8]);
    let mut b = a.clone();
    b += &a;
    b *= 2;
    assert_eq!(b.to_vec(), vec![4, 8, 12, 16]);
    b -= &a;
    b /= 3;
    assert_eq!(b.to_vec(), vec![1, 2, 3, 4]);
}

#[test]
fn vec_eq_and_clone() {
    let a = Vector::<TestType>::from_vec(vec![1, 2, 3, 4]);
    let b = Vector::<TestType>::from_vec(vec![1, 2, 3, 5]);
    let c = Vector::<TestType>::from_vec(vec![1, 2, 3]);
    let d = Vector::<TestType>::from_vec(vec![1, 2, 3, 4]);
    assert!(a == d && a != b && a != c && c != a);
    assert!(a.clone() == a);
}

#[test]
fn mat_eq() {
    let a = Matrix::<TestType>::from_vec((0..12).collect(), 4, 3);
    let mut other: Vec<TestType> = (0..12).collect();
    other[11] = 12;
    let b = Matrix::from_vec(other, 4, 3);
    let c = Matrix::<TestType>::from_vec((0..12).collect(), 3, 4);
    assert!(a == a.clone() && a != b && a != c);
}

#[test]
fn length_and_normalize() {
    for &size in &[1usize, 2, 8, 16, 64, 128] {
        let v: Vec<f32> = (0..size).map(|x| x as f32).collect();
        let expect = v.iter().map(|x| x * x).sum::<f32>().sqrt();
        assert!((Vector::from_vec(v).length() - expect).abs() < 1e-4);
        let w: Vec<f32> = (1..=size).map(|x| x as f32).collect();
        assert!((Vector::from_vec(w).normalize().length() - 1.0).abs() < 0.01);
    }
}

#[test]
fn save_load() {
    let dir = std::env::temp_dir();
    let vp = dir.join("linalg_b200_vec.bin");
    let a = Vector::<TestType>::from_vec(vec![19088743, 2309737967, 4275878552, 1985229328]);
    a.save(vp.to_str().unwrap()).unwrap();
    assert!(Vector::<TestType>::open(vp.to_str().unwrap()).unwrap() == a);
    let mp = dir.join("linalg_b200_mat.bin");
    let m = Matrix::<TestType>::from_vec(vec![19088743, 2309737967, 324508639, 4275878552, 1985229328, 38177486], 2, 3);
    m.save(mp.to_str().unwrap()).unwrap();
    assert!(Matrix::<TestType>::open(mp.to_str().unwrap()).unwrap() == m);
    assert!(Matrix::<u16>::open(mp.to_str().unwrap()).is_err()); // element size mismatch
}

#[test]
fn outer_product_and_transpose() {
    let o = mul_column_row(&Vector::<TestType>::from_vec(vec![1, 2]), &Vector::from_vec(vec![3, 4, 5]));
    assert!(o == Matrix::from_vec(vec![3, 4, 5, 6, 8, 10], 2, 3));
    assert!(o.transpose() == Matrix::from_vec(vec![3, 6, 4, 8, 5, 10], 3, 2));
}

#[test]
fn user_kernels() {
    let src = "kernel void {T}_twice(global {T}* C, global {T}* B) { C[i] = B[i] * ({T})(2); }\n\
               kernel void {T}_blend(global {T}* C, global {T}* A, global {T}* B) { C[i] = A[i] * B[i] + A[i]; }\n\
               kernel void {T}_ramp(global {T}* C, {T} first) { C[i] = first + ({T})i; }\n";
    let ty = <TestType as Parameter>::type_to_str();
    let mut twice = create_kernel_from_source::<TestType>(&format!("{}_twice", ty), src).arg_buf_named("C").arg_buf_named("B");
    let mut blend = create_kernel_from_source::<TestType>(&format!("{}_blend", ty), src).arg_buf_named("C").arg_buf_named("A").arg_buf_named("B");
    let mut ramp = create_kernel_from_source::<TestType>(&format!("{}_ramp", ty), src).arg_buf_named("C").arg_scl_named("first", 10 as TestType);
    let a = Vector::<TestType>::from_vec(vec![1, 2, 3, 4, 5]);
    let b = Vector::<TestType>::from_vec(vec![5, 4, 3, 2, 1]);
    assert_eq!(a.map(&mut twice).to_vec(), vec![2, 4, 6, 8, 10]);
    assert_eq!(Vector::from_for_each2(&a, &b, &mut blend).to_vec(), vec![6, 10, 12, 12, 10]);
    let mut c = a.clone();
    c.for_each_mut(&a, &mut twice);
    assert_eq!(c.to_vec(), vec![2, 4, 6, 8, 10]);
    let g = Vector::<TestType>::generate(&mut ramp, 300);
    assert_eq!(g.to_vec(), (10..310).collect::<Vec<TestType>>());
}

#[test]
#[should_panic]
fn shape_mismatch_panics() {
    let a = Vector::<TestType>::from_vec(vec![1, 2, 3]);
    let b = Vector::<TestType>::from_vec(vec![1, 2]);
    let _ = &a + &b;
}
